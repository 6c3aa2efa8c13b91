"""Child process of test_null_arguments_are_refused: every C-ABI entry point
that can be reached without a GPU is called with NULL / out-of-range arguments.
Each must return AMRVR_ERR_INVALID (1) -- or be a no-op for the free/close
family -- and none may dereference anything.  Prints "ok <n>" at the end."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

from amrvolumerenderer_b200 import ctypes_api as api  # noqa: E402

lib = api.load_library()
cam = api.Camera()
i, j = C.c_int(), C.c_int()
one, zero = C.c_float(1.0), C.c_float(0.0)
table = (C.c_float * 1024)()
N = None

REFUSED = [  # (name, call) -> status 1
    ("order_boxes_host null table", lambda: lib.amrvr_order_boxes_host(2, 3, N, C.byref(cam), one, 1, N, N)),
    ("order_boxes_host negative count", lambda: lib.amrvr_order_boxes_host(2, -1, N, C.byref(cam), one, 1, N, N)),
    ("order_boxes_host no ranks", lambda: lib.amrvr_order_boxes_host(0, 0, N, C.byref(cam), one, 1, N, N)),
    ("layer_order_host", lambda: lib.amrvr_layer_order_host(3, N, N, N, N)),
    ("piece_range no pieces", lambda: lib.amrvr_piece_range(10, 0, 0, C.byref(i), C.byref(j))),
    ("piece_range null out", lambda: lib.amrvr_piece_range(10, 0, 2, N, N)),
    ("piece_range negative", lambda: lib.amrvr_piece_range(-1, 5, 2, C.byref(i), C.byref(j))),
    ("reference_sample_distance", lambda: lib.amrvr_reference_sample_distance(2, N, N, N, N)),
    ("color_table null", lambda: lib.amrvr_color_table(one, one, zero, one, N, 3, N)),
    ("color_table negative count", lambda: lib.amrvr_color_table(one, one, zero, one, N, -3, table)),
    ("auto_camera", lambda: lib.amrvr_auto_camera(N, N, 0, N, N)),
    ("exchange_plan_host", lambda: lib.amrvr_exchange_plan_host(2, 0, 3, N, N, 8, 8, N, N, N, N, N)),
    ("convexify_boxes", lambda: lib.amrvr_convexify_boxes(2, N, 0, N, N, N, N, 0, N)),
    ("convexify_levels", lambda: lib.amrvr_convexify_levels(2, N, N, N, N, N, 0, N)),
    ("plotfile_open", lambda: lib.amrvr_plotfile_open(N, N)),
    ("plotfile_get_info", lambda: lib.amrvr_plotfile_get_info(N, N)),
    ("plotfile_get_level", lambda: lib.amrvr_plotfile_get_level(N, 0, N)),
    ("plotfile_level_boxes", lambda: lib.amrvr_plotfile_level_boxes(N, 0, N, 0)),
    ("plotfile_read_fab", lambda: lib.amrvr_plotfile_read_fab(N, 0, 0, 0, N, 0)),
    ("plotfile_read_level", lambda: lib.amrvr_plotfile_read_level(N, 0, 0, 0, 1, 0, 1, N, 0, N, N, 0, N)),
    ("ctx_create null out", lambda: lib.amrvr_ctx_create(0, 0, 1, N, N)),
    ("scene_upload null ctx", lambda: lib.amrvr_scene_upload(N, 1, N, N, zero, one, N)),
    ("synchronize null ctx", lambda: lib.amrvr_synchronize(N)),
]
NO_OPS = [  # free/close/query family on NULL handles: must simply return
    lambda: lib.amrvr_plotfile_close(N),
    lambda: lib.amrvr_ctx_destroy(N),
    lambda: lib.amrvr_scene_free(N, N),
    lambda: lib.amrvr_layers_free(N, N),
    lambda: lib.amrvr_plotfile_var_name(N, 0),
    lambda: lib.amrvr_ctx_kernel_launches(N),
]

for name, call in REFUSED:
    status = call()
    assert status == 1, f"{name}: status {status}"
    assert lib.amrvr_last_error(None), name
assert lib.amrvr_plotfile_var_index(N, b"x") == -1
for call in NO_OPS:
    call()
print("ok", len(REFUSED) + len(NO_OPS) + 1)
