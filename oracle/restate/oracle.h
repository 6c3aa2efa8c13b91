/* oracle/restate/oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of amrVolumeRenderer's hot path (ray-march of AMR
 * bricks -> per-box layers -> visibility ordering -> sort-last DirectSend
 * composite -> downsample / overlay / byte quantisation).  Every function cites
 * the reference file:line it follows.  Only tests/, __graft_entry__.smoke() and
 * bench.py's CPU-baseline legs may load this library; the product
 * (amrvolumerenderer_b200/) never does.
 *
 * Pinning: the reference's own tests hold no golden vector for this path
 * (SURVEY.md section 4), so the restatement is pinned against the reference's
 * own sources compiled here (oracle/_ref, see oracle/Makefile) by
 * tests/test_oracle_vs_ref.py, and against fixtures generated from oracle/_ref
 * committed under tests/golden/.
 *
 * Arithmetic rules: compiled WITHOUT FMA contraction (-ffp-contract=off, no
 * -march), float expressions stay float (x86-64 SSE2), libm calls use the same
 * overloads the C++ reference resolves to (powf/tanf/sqrtf for float
 * arguments, pow/log for double).
 */
#ifndef AMRVR_ORACLE_H
#define AMRVR_ORACLE_H

#include "../../include/amrvr.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Common/VolumePainter.cpp:442-516 buildColorTable -> 256 RGBA floats. */
void orc_color_table(float alpha_scale, float normalization_factor,
                     float range_min, float range_max,
                     const amrvr_colormap_pt* pts, int npts, float* table);

/* Common/VolumePainter.cpp:571-610: per-box step and normalisation factor. */
void orc_box_sampling(const amrvr_box* box, const double* bounds_min,
                      const double* bounds_max, float reference_sample_distance,
                      float* sample_distance, float* normalization_factor);

/* Common/VolumeTypes.hpp:33-67 + VolumePainter.cpp:873-883: colour-table index
 * of one raw cell value. */
int orc_cell_index(double raw, const amrvr_scalar_transform* t,
                   float range_min, float range_max);
/* The same for every cell of a box, i fastest. */
void orc_quantize_box(const amrvr_box* box, const amrvr_scalar_transform* t,
                      float range_min, float range_max, uint8_t* out);

/* Common/VolumePainter.cpp:548-961 VolumePainter::paint -> dense layer
 * (width*height*5 floats).  samples (may be NULL) counts executions of the
 * body at :869-894. */
void orc_paint(const amrvr_box* box, const amrvr_scalar_transform* t,
               float range_min, float range_max, float box_transparency,
               float reference_sample_distance, int width, int height,
               const amrvr_camera* camera, const amrvr_colormap_pt* pts,
               int npts, float* out5, long long* samples);

/* VolumeRenderer/VolumeRenderer.cpp:541-553 computeBoxDepthHint. */
float orc_depth_hint(const amrvr_box_record* rec, const amrvr_camera* camera);
/* DirectSend/Base/DirectSendBase.cpp:363-388: order[k] = global box index. */
void orc_layer_order(int nglobal, const amrvr_box_record* recs,
                     const amrvr_camera* camera, float* hints, int* order);
/* VolumeRenderer.cpp:1138-1190 referenceSampleDistance. */
float orc_reference_sample_distance(int nglobal, const amrvr_box_record* recs,
                                    const double* bounds_min,
                                    const double* bounds_max);

/* Common/VisibilityOrdering.cpp:63-632.  Returns 1 when the graph ordering
 * succeeded, 0 when the identity fallback was taken. */
int orc_order(int nranks, int nglobal, const amrvr_box_record* recs,
              const amrvr_camera* camera, float aspect, int use_graph,
              int* rank_order);

/* DirectSend/Base/DirectSendBase.cpp:59-74 getPieceRange. */
void orc_piece_range(int num_pixels, int piece, int num_pieces, int* begin,
                     int* end);
/* Common/ImageRGBAFloatColorDepthSort.hpp:13-27 Features::blend. */
void orc_blend(const float* top, const float* bottom, float* out);
/* DirectSendBase.cpp:316-458 composeLayered + ImageColorOnly.hpp:220-270 Gather:
 * dense layers in, gathered frame (width*height*5) and per-pixel owning rank
 * out.  layer l belongs to rank layer_owner[l]; layers of a rank keep their
 * relative order. */
void orc_compose(int nranks, int width, int height, int nlayers,
                 const int* layer_owner, const float* const* layer_data,
                 const float* depth_hints, const int* rank_order, float* full5,
                 int* piece_owner);

/* VolumeRenderer.cpp:479-528 downsampleImage. */
void orc_downsample(const float* src5, int width, int height, int sqrt_aa,
                    float* dst5);
/* VolumeRenderer.cpp:139-335 renderBoundingBoxLayer (sqrtAA = 1 at :1313). */
void orc_overlay(const double* tight_min, const double* tight_max,
                 const amrvr_camera* camera, int sqrt_aa, int width, int height,
                 float* image5);
/* Common/Color.hpp:66-91 + Common/SavePPM.cpp:25-32: RGB8, top row first. */
void orc_quantize_rgb8(const float* image5, int width, int height,
                       uint8_t* rgb8);

/* VolumeRenderer.cpp:1103-1339 renderSingleTrial on `nranks` emulated ranks:
 * boxes[i] is painted by owners[i].  Outputs (any may be NULL): final5
 * (params->width*height*5 after downsample + overlay), rgb8, rank_order,
 * samples. */
int orc_render_frame(int nranks, int nboxes, const amrvr_box* boxes,
                     const int32_t* owners, const amrvr_scalar_transform* t,
                     float range_min, float range_max, const double* bounds_min,
                     const double* bounds_max, const amrvr_camera* camera,
                     const amrvr_frame_params* params,
                     const amrvr_colormap_pt* pts, int npts, float* final5,
                     uint8_t* rgb8, int* rank_order, long long* samples);

/* VolumeRenderer.cpp:974-1023: the automatic camera for explicit bounds.
 * Takes the two uniform_real_distribution<float> draws as inputs (libstdc++'s
 * mt19937/uniform_real_distribution are not restated in C). */
void orc_auto_camera(const double* bounds_min, const double* bounds_max,
                     float azimuth, float altitude, const double* up_or_null,
                     amrvr_camera* out);

/* VolumeRenderer/SceneBuilder.cpp:113-425 BuildSceneGeometry for already-convex
 * levels holding the FABs of ALL ranks (orc_scene.c).  Returns 0, or 1 "no
 * data", 2 "invalid scalar range", 3 no positive value for the log scale,
 * 4 normalisation span not finite, 5 boxes_out too small. */
int orc_scene_build(int nlevels, const amrvr_level* levels, int min_level,
                    int max_level, int component, int log_scale_input,
                    int normalize_to_data_range, amrvr_box* boxes_out,
                    int capacity, amrvr_scene_summary* summary);

/* VolumeRenderer/SceneBuilder.cpp:445-577 ComputeSceneHistogram. */
void orc_scene_histogram(int nboxes, const amrvr_box* boxes,
                         const amrvr_scalar_transform* t, float range_min,
                         float range_max, int bin_count, uint64_t* counts,
                         uint64_t* sample_count);

#ifdef __cplusplus
}
#endif
#endif
