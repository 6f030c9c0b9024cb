// kernels.cuh — launchers of the wavefront kernels (definitions in kernels.cu).
#pragma once
#include "../../include/pt_b200.h"
#include "device_scene.cuh"

namespace ptb {
// fixed ray batches through the wave queues (the parity harness of k_extend / k_shadow)
void launch_load_rays(const float *rays, uint32_t n, uint32_t slot_base, const WaveState &W, uint32_t buf, cudaStream_t st);
void launch_load_shadow(const float *rays, const float *limits, const float *prim, uint32_t n, const WaveState &W,
                        uint32_t which, cudaStream_t st);
void launch_store_transmittance(const WaveState &W, uint32_t n, float *out, cudaStream_t st);
void launch_trace_paths(const DevScene &S, const WaveParams &P, const uint32_t *pixels, uint32_t n, pt_path_vertex *out,
                        uint32_t *counts, cudaStream_t st);
void launch_philox(const uint32_t *ctr, const uint32_t *key, uint32_t n, uint32_t *out, cudaStream_t st);
void launch_triangle_intersect(const float *rays, const float *tris, uint32_t n, pt_hit *out, cudaStream_t st);
void launch_debug_aov(const DevScene &S, uint32_t width, uint32_t height, uint8_t *planes, uint32_t *n_hit, cudaStream_t st);
void launch_init_srgb_lut(float *lut, cudaStream_t st);
void launch_raygen(const DevScene &S, const WaveParams &P, const WaveState &W, cudaStream_t st);
// alpha = false: first element of ray_cast's list; trace_out (optional): pt_hit per queue position
void launch_extend(const DevScene &S, const WaveParams &P, const WaveState &W, uint32_t bounce, bool alpha,
                   pt_hit *trace_out, cudaStream_t st);
void launch_hit_list(const DevScene &S, const WaveParams &P, const WaveState &W, uint32_t bounce, cudaStream_t st);
void launch_shade(const DevScene &S, const WaveParams &P, const WaveState &W, uint32_t bounce, cudaStream_t st);
int shade_launches();  // launches of launch_hit_list + launch_shade together
void launch_shadow(const DevScene &S, const WaveState &W, uint32_t bounce, cudaStream_t st);
void launch_resolve(const WaveParams &P, const WaveState &W, float *accum, cudaStream_t st);
void launch_accum_add(float *dst, const float *src, uint32_t n, cudaStream_t st);
void launch_tonemap(const float *accum, uint32_t npix, uint32_t total_samples, int tonemap, uint8_t *rgb8, cudaStream_t st);
// pt_trace_closest_device_noalloc: closest hits of n <= PT_TRACE_NOALLOC_MAX_RAYS rays with the wave scratch in a
// static device array (kernels.cu); allocates nothing
cudaError_t launch_trace_closest_static(const DevScene &S, const float *rays_dev, uint32_t n, pt_hit *out_dev, cudaStream_t st);
}  // namespace ptb
