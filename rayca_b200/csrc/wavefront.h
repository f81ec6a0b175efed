// wavefront.h — wavefront scheduling with ray sorting for the path tracer's bounce rays.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>

#include <cuda_runtime.h>

namespace rb {

struct RenderParams;

// device scratch of the wavefront path, owned by the context, grown on demand and reused across frames
struct WaveScratch {
    void* base = nullptr;
    size_t bytes = 0;
};

// true when this launch can run as wavefront stages: Pathtracer with at least one bounce, cosine sampler, one light
// sample, one sample per pixel, point lights, the product walk (traversal mode 0 / 3), the static grid
// ... and enough paths for the extra launches to pay (`rows` x width of this launch)
bool wave_supported(const RenderParams& P, uint32_t rows, uint32_t mode, bool count, bool persistent, bool general_integrator, bool bounce);

// Enqueues the whole frame on `st`: stage 0 (primary rays, depth 0 exactly as the fused kernel), then per bounce depth
// a radix sort of the live rays by (direction octant, Morton code of the origin) and a stage that traces them in that
// order.  Returns RAYCA_OK or an error code; `launches` receives the number of kernels of this library launched
// (the sort's own kernels not counted).
int wave_render(const RenderParams& P, uint32_t rows, bool spheres, WaveScratch& ws, cudaStream_t st, uint32_t* launches, std::string& err);

} // namespace rb
