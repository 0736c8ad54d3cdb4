// launch.hpp — host-callable launchers exported by the per-dtype translation units.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace dlnb {
struct EmuViews;
cudaError_t launch_coll_f32(int op, int nr, dim3 grid, int threads, cudaStream_t s,
                            const EmuViews &ev, size_t count, bool coop, bool pdl);
cudaError_t launch_coll_f16(int op, int nr, dim3 grid, int threads, cudaStream_t s,
                            const EmuViews &ev, size_t count, bool coop, bool pdl);
cudaError_t launch_coll_bf16(int op, int nr, dim3 grid, int threads, cudaStream_t s,
                             const EmuViews &ev, size_t count, bool coop, bool pdl);
int occupancy_coll_f32(int op, int nr, int threads);
int occupancy_coll_f16(int op, int nr, int threads);
int occupancy_coll_bf16(int op, int nr, int threads);
} // namespace dlnb
