/* TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * pt_oracle.h -- plain-C CPU restatement of the reference's per-pixel path-tracing function
 * (src/raytracer.cu:116-136 and everything it calls).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline leg may use it, and only as the checker.
 * Parity status: PINNED against oracle/_ref/libref_host.so (the reference's own source compiled
 * for the host) in this container, and against the golden fixtures under tests/golden/ that were
 * generated from it (tests/golden/make_golden.py).
 */
#ifndef PT_ORACLE_H
#define PT_ORACLE_H
#include "../include/rtb.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct PtoScene PtoScene;

PtoScene *pto_scene_create(const RtbSceneDesc *desc);
void pto_scene_destroy(PtoScene *s);

/* one frame, rows [y0, y1): out/prev are full width*height*3 images (prev may be NULL = zeros) */
int pto_render(const PtoScene *s, const RtbCamData *cam, int width, int height, const RtbRenderData *rd,
               int time_ms, int frame_num, const float *prev, float *out, int y0, int y1, int nthreads,
               RtbCounters *counters /* nullable */);
int pto_primary_ids(const PtoScene *s, const RtbCamData *cam, int width, int height, int32_t *obj_id, int32_t *sub_id,
                    float *t, int nthreads);

/* parse_pixel_colours, src/main.cu:343-371: float RGB -> RGBA8 (truncation, clamp to [0, 255], alpha 255) */
void pto_display_rgba8(const float *rgb, int width, int height, unsigned char *rgba);

/* known-answer hooks */
unsigned int pto_seed(int array_index, int time_ms);
float pto_rng(unsigned int *state);
float pto_normal(unsigned int *state);
void pto_primary_dir(const RtbCamData *cam, int x, int y, float *dir3);
/* Camera::assign_constant_mem at the default pose with run-time W/H (src/camera.cu:46-108) */
void pto_default_camera(int width, int height, RtbCamData *out);

#ifdef __cplusplus
}
#endif
#endif
