"""B200-native hot path of amrVolumeRenderer behind the C ABI of include/amrvr.h.

The reference's Python package exports render, compute_histogram,
initialize_runtime and finalize_runtime (python/amrVolumeRenderer/__init__.py);
here they live in the `render` submodule with the same keywords:

    from amrvolumerenderer_b200.render import (render, compute_histogram,
                                               initialize_runtime, finalize_runtime)

(`render` is also the submodule's name, so the package itself re-exports only
the other three, resolved on first use to keep submodule imports cheap.)"""

__all__ = ["initialize_runtime", "finalize_runtime", "compute_histogram"]


def __getattr__(name):
    if name in __all__:
        import importlib

        return getattr(importlib.import_module(".render", __name__), name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
