// dlap_panel_image: the X images of every tile of the valid-cell index, made once per panel.
// For tile k (<= 128 listed stocks of one month) the image pair holds what the gather warps of the tensor-core
// kernels would put into shared memory for that tile (cell_tc.cuh: no-swizzle canonical layout [6][128+1][8] bf16,
// hi image then lo image; the ones column at k = 46; rows past the end of the tile repeat its last row), so a step
// can fetch a tile with one bulk copy instead of gathering and splitting 128 rows of x, twice per step.
// 2 * X_IMG = 24,768 bytes per tile at [k * 24768); HBM-bound byte work, once per panel.
#include "cell_tc.cuh"

namespace dlap {
using namespace tcl;

namespace {
constexpr int IMG_T = 128;     // 4 warps, 32 rows each

template <bool WIDE>
__device__ __forceinline__ void image_tiles(const CellParams& P, uint8_t* img, uint8_t* sm) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles = __ldg(P.index);
  init_x_images(sm, sm + X_IMG, tid, IMG_T);            // ones column, zero padding: never overwritten below
  __syncthreads();
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    MonthCursor mc;
    mc.init(P, tile);
    const int cell = warp * 32 + lane;
    const int n = __ldg(vidx_row(P, mc.t) + mc.j0(tile) + min(cell, mc.ncell(tile) - 1));
    uint64_t rows[32];
    gather_load<32, WIDE>(P, mc.t, n, lane, rows);
    rows_to_image<32>(rows, sm, sm + X_IMG, warp * 32, lane, P.F);
    __syncthreads();
    uint4* dst = reinterpret_cast<uint4*>(img + (size_t)tile * (2 * X_IMG));
    const uint4* src = reinterpret_cast<const uint4*>(sm);
    for (int i = tid; i < (int)(2 * X_IMG / 16); i += IMG_T) dst[i] = src[i];
    __syncthreads();
  }
}

__global__ void __launch_bounds__(IMG_T) panel_image_kernel(const CellParams P, uint8_t* img) {
  extern __shared__ __align__(16) uint8_t sm_img[];
  if (gather_wide(P)) image_tiles<true>(P, img, sm_img);
  else image_tiles<false>(P, img, sm_img);
}
}  // namespace

size_t panel_image_tile_bytes() { return 2 * X_IMG; }

cudaError_t launch_panel_image(const CellParams& P, uint8_t* image, int sms, cudaStream_t st) {
  const long long bound = (long long)P.T * ((P.N + TILE - 1) / TILE);
  const int grid = (int)(bound < (long long)sms * 8 ? bound : (long long)sms * 8);
  panel_image_kernel<<<grid, IMG_T, 2 * X_IMG, st>>>(P, image);
  return cudaGetLastError();
}

}  // namespace dlap
