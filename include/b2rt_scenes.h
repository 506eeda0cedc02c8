/*
 * b2rt_scenes.h -- C ABI of libb2rt_scenes.so: seeded procedural stand-ins for the scenes the reference is quoted on.
 * Benchmark / test infrastructure, host only, no CUDA: NOT part of the drop-in boundary (include/b2rt.h) and not
 * linked into libb2rt.so.  Both arms of bench.py get their geometry from here (the reference arm writes it out as
 * an OBJ for the reference's own importer), so neither needs the other's library.
 */
#ifndef B2RT_SCENES_H
#define B2RT_SCENES_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Seeded procedural stand-ins for the README scenes (no network: bunny/conference/dragon/
 * hairball cannot be fetched).  kind: 0 blob (bunny-scale), 1 interior (conference-scale),
 * 2 knot (dragon-scale), 3 strands (hairball-scale).  SURVEY.md 8(d). */
typedef struct b2rt_scene b2rt_scene;      /* every call returns 0 on success, -1 on a bad argument */
int  b2rt_scene_generate(int kind, int target_tris, uint64_t seed, b2rt_scene** out);
void b2rt_scene_free(b2rt_scene* s);
int  b2rt_scene_arrays(const b2rt_scene* s, const float** verts, int* num_verts,
                       const int32_t** tris, int* num_tris);
/* camera for the scene: position[3], forward[3], up[3], fov, near, far (12 floats) */
int  b2rt_scene_camera(const b2rt_scene* s, float cam[12]);
int  b2rt_scene_write_obj(const b2rt_scene* s, const char* path);

#ifdef __cplusplus
}
#endif
#endif /* B2RT_SCENES_H */
