// TMA probe: load one {Wt, 1, 32} fp32 box of a [NC, H, W] tensor with a given swizzle / L2 promotion and dump the
// shared-memory image, so that the layout formulas used by roi_align.cu can be checked on the device.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/tma_probe tools/tma_probe.cu -lcuda
//   tools/tma_probe <swizzle 0..3> <Wt> <l2promo 0..3> [rows_per_box]
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

__global__ void probe(const __grid_constant__ CUtensorMap map, float* out, int nfloats, int x, int y, int c, unsigned bytes) {
  extern __shared__ unsigned char raw[];
  unsigned base = (unsigned)__cvta_generic_to_shared(raw);
  unsigned al = (base + 1023u) & ~1023u;
  float* buf = reinterpret_cast<float*>(raw + (al - base));
  unsigned bar = al + 16384;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(al),
                 "l"(&map), "r"(bar), "r"(x), "r"(y), "r"(c)
                 : "memory");
  }
  unsigned done = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar) : "memory");
  }
  for (int i = threadIdx.x; i < nfloats; i += blockDim.x) out[i] = buf[i];
}

int main(int argc, char** argv) {
  int swz = argc > 1 ? atoi(argv[1]) : 0, Wt = argc > 2 ? atoi(argv[2]) : 8, promo = argc > 3 ? atoi(argv[3]) : 0;
  int rows = argc > 4 ? atoi(argv[4]) : 1;
  int X = argc > 5 ? atoi(argv[5]) : 5;
  const int W = 64, H = 16, NC = 64;
  std::vector<float> h((size_t)W * H * NC);
  for (int c = 0; c < NC; ++c)
    for (int y = 0; y < H; ++y)
      for (int x = 0; x < W; ++x) h[((size_t)c * H + y) * W + x] = c * 10000 + y * 100 + x;   // value encodes (c, y, x)
  float *d, *o;
  cudaMalloc(&d, h.size() * 4);
  cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
  const int nfloats = Wt * rows * 32;
  cudaMalloc(&o, nfloats * 4);
  CUtensorMap map;
  cuuint64_t dims[3] = {W, H, NC};
  cuuint64_t strides[2] = {W * 4, (cuuint64_t)W * H * 4};
  cuuint32_t box[3] = {(cuuint32_t)Wt, (cuuint32_t)rows, 32};
  cuuint32_t es[3] = {1, 1, 1};
  cuInit(0);
  CUresult r = cuTensorMapEncodeTiled(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, d, dims, strides, box, es,
                                      CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)swz,
                                      (CUtensorMapL2promotion)promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("swz=%d Wt=%d promo=%d rows=%d: encode failed %d\n", swz, Wt, promo, rows, (int)r); return 0; }
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
  probe<<<1, 128, 32768>>>(map, o, nfloats, X, 3, 32, (unsigned)nfloats * 4);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("swz=%d Wt=%d promo=%d rows=%d: kernel failed: %s\n", swz, Wt, promo, rows, cudaGetErrorString(e)); return 0; }
  std::vector<float> res(nfloats);
  cudaMemcpy(res.data(), o, nfloats * 4, cudaMemcpyDeviceToHost);
  // check the layout formula: element (c, x) of the row at float offset c*Wt + 4*((x/4) ^ sw(c)) + x%4
  int bad = 0;
  for (int rr = 0; rr < rows; ++rr)
  for (int c = 0; c < 32; ++c)
    for (int x = 0; x < Wt; ++x) {
      int rowidx = c * rows + rr;   // smem rows are (c-major, y-minor)
      int sw = swz == 3 ? (rowidx & 7) : swz == 2 ? ((rowidx >> 1) & 3) : swz == 1 ? ((rowidx >> 2) & 1) : 0;
      float got = res[(size_t)rowidx * Wt + 4 * ((x / 4) ^ sw) + x % 4];
      float want = (32 + c) * 10000 + (3 + rr) * 100 + (X + x);
      if (got != want) ++bad;
    }
  printf("swz=%d Wt=%d promo=%d rows=%d: ok, layout mismatches %d  first floats %g %g %g %g | %g\n", swz, Wt, promo, rows, bad, res[0], res[1],
         res[4], res[Wt], res[nfloats - 1]);
  return 0;
}
