// tma_inorder.cu - does the data of a CTA's FIRST tile arrive early when the CTA issues the TMA copies of all its tiles
// in tile order?  One CTA per SM slot (grid = SMs x CTAS), each owns tiles blockIdx.x, +grid, ... of a 16 384-env
// BlindGrasping-shaped input (L2-cold ring); one thread issues every copy of every tile back to back, one mbarrier per tile;
// a second thread records the globaltimer at which each tile's barrier completes.  Also: the same with 4-byte LDG gathers
// issued tile by tile by 8 warps (register staged, the shipped single-pass pattern) for comparison.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_inorder tools/microbench/tma_inorder.cu && ./tma_inorder
#include <cuda.h>
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int NB = 32, NA = 2, KST = 85, MAXT = 8;
__constant__ int c_rb[11] = {0, 5, 11, 17, 23, 29, 6, 12, 18, 24, 30};
__constant__ int c_cf[6] = {4, 10, 16, 22, 28, 31};
struct Bufs { const float *rb, *cf, *root, *dof, *st; float* out; unsigned long long* t_land; int n, tiles; };
struct Maps { CUtensorMap rb16, rb12, cf8, root16, st; };

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ void tma2d(uint32_t dst, const CUtensorMap* m, int c0, int c1, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               :: "r"(dst), "l"(m), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
constexpr uint32_t O_DOF = 0, O_HAND = 6656, O_TP = O_HAND + 2048, O_CF = O_TP + 10 * 1536, O_ROOT = O_CF + 6 * 1024,
                   O_ST = O_ROOT + 2 * 1024, TILE_BYTES = O_ST + KST * 128;  // 45 952 B per tile buffer

// all tiles of the CTA are resident in shared memory at once (<= 4 tiles x 45 KB)
template <int ISSUERS>
__global__ void __launch_bounds__(256, 1) k_tma(Bufs b, const __grid_constant__ Maps M, unsigned long long* t0_out) {
  extern __shared__ __align__(128) unsigned char sm[];
  __shared__ __align__(8) unsigned long long bar[MAXT];
  const int tid = threadIdx.x;
  int nt = 0;
  for (int t = blockIdx.x; t < b.tiles; t += gridDim.x) ++nt;
  if (tid == 0) {
    for (int k = 0; k < nt; ++k) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(s32(&bar[k])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (blockIdx.x == 0) *t0_out = gtime();
  }
  __syncthreads();
  if (tid == 0)
    for (int k = 0; k < nt; ++k)
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(s32(&bar[k])), "r"(TILE_BYTES) : "memory");
  __syncthreads();
  if (tid < ISSUERS) {
    for (int k = 0; k < nt; ++k) {
      const int e0 = (blockIdx.x + k * gridDim.x) * 32;
      const uint32_t base = s32(sm) + k * TILE_BYTES, mb = s32(&bar[k]);
      for (int c = tid; c < 21; c += ISSUERS) {
        if (c == 0)
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       :: "r"(base + O_DOF), "l"(b.dof + (size_t)e0 * 52), "r"(6656u), "r"(mb) : "memory");
        else if (c == 1) tma2d(base + O_HAND, &M.rb16, 0, e0, mb);
        else if (c < 12) tma2d(base + O_TP + (c - 2) * 1536, &M.rb12, (c_rb[c - 1] * 13) & ~3, e0, mb);
        else if (c < 18) tma2d(base + O_CF + (c - 12) * 1024, &M.cf8, (c_cf[c - 12] * 3) & ~3, e0, mb);
        else if (c < 20) tma2d(base + O_ROOT + (c - 18) * 1024, &M.root16, (13 + 26 * (c - 18)) & ~3, e0 / 2, mb);
        else tma2d(base + O_ST, &M.st, e0, 0, mb);
      }
    }
  }
  if (tid == 32) {  // the observer thread: landing time of every tile, in order
    for (int k = 0; k < nt; ++k) {
      uint32_t ok = 0;
      while (!ok)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0, 1000000; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(s32(&bar[k])) : "memory");
      b.t_land[(size_t)blockIdx.x * MAXT + k] = gtime();
    }
  }
  __syncthreads();
  float acc = 0.f;
  for (uint32_t i = tid; i < nt * TILE_BYTES / 4; i += 256) acc += reinterpret_cast<float*>(sm)[i];
  b.out[(size_t)blockIdx.x * 256 + tid] = acc;
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static CUtensorMap make2d(EncodeFn enc, const float* base, uint64_t inner, uint64_t outer, uint64_t pitch, uint32_t b0, uint32_t b1) {
  CUtensorMap m;
  cuuint64_t dims[2] = {inner, outer}, strides[1] = {pitch};
  cuuint32_t box[2] = {b0, b1}, es[2] = {1, 1};
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
  return m;
}

int main(int argc, char** argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 16384, ring = 6, tiles = n / 32;
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  EncodeFn enc = (EncodeFn)fn;
  const int grid = sms;
  if ((tiles + grid - 1) / grid > 4) { printf("too many tiles per CTA for this microbenchmark\n"); return 1; }
  std::vector<Bufs> bufs(ring);
  std::vector<Maps> maps(ring);
  unsigned long long* t0;
  CK(cudaMalloc(&t0, 8));
  for (int i = 0; i < ring; ++i) {
    float *rb, *cf, *root, *dof, *st, *out;
    unsigned long long* tl;
    CK(cudaMalloc(&rb, (size_t)n * NB * 13 * 4)); CK(cudaMalloc(&cf, (size_t)n * NB * 3 * 4)); CK(cudaMalloc(&root, (size_t)n * NA * 13 * 4));
    CK(cudaMalloc(&dof, (size_t)n * 52 * 4)); CK(cudaMalloc(&st, (size_t)KST * n * 4)); CK(cudaMalloc(&out, (size_t)grid * 256 * 4));
    CK(cudaMalloc(&tl, (size_t)grid * MAXT * 8));
    CK(cudaMemset(rb, 0, (size_t)n * NB * 13 * 4)); CK(cudaMemset(cf, 0, (size_t)n * NB * 3 * 4)); CK(cudaMemset(root, 0, (size_t)n * NA * 13 * 4));
    CK(cudaMemset(dof, 0, (size_t)n * 52 * 4)); CK(cudaMemset(st, 0, (size_t)KST * n * 4));
    bufs[i] = Bufs{rb, cf, root, dof, st, out, tl, n, tiles};
    maps[i].rb16 = make2d(enc, rb, NB * 13, n, NB * 13 * 4, 16, 32);
    maps[i].rb12 = make2d(enc, rb, NB * 13, n, NB * 13 * 4, 12, 32);
    maps[i].cf8 = make2d(enc, cf, NB * 3, n, NB * 3 * 4, 8, 32);
    maps[i].root16 = make2d(enc, root, NA * 13 * 2, n / 2, NA * 13 * 2 * 4, 16, 16);
    maps[i].st = make2d(enc, st, n, KST, (uint64_t)n * 4, 32, KST);
  }
  const size_t smem = 4 * TILE_BYTES;
  CK(cudaFuncSetAttribute(k_tma<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(k_tma<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for (int variant = 0; variant < 2; ++variant) {
    for (int warm = 0; warm < 3; ++warm)
      for (int i = 0; i < ring; ++i) {
        if (variant == 0) k_tma<1><<<grid, 256, smem>>>(bufs[i], maps[i], t0);
        else k_tma<4><<<grid, 256, smem>>>(bufs[i], maps[i], t0);
      }
    CK(cudaDeviceSynchronize());
    // timing
    const int reps = 50;
    CK(cudaEventRecord(e0));
    for (int r = 0; r < reps; ++r)
      for (int i = 0; i < ring; ++i) {
        if (variant == 0) k_tma<1><<<grid, 256, smem>>>(bufs[i], maps[i], t0);
        else k_tma<4><<<grid, 256, smem>>>(bufs[i], maps[i], t0);
      }
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    // landing times of the last launch on replica 0 (cold: 5 other replicas ran since)
    if (variant == 0) k_tma<1><<<grid, 256, smem>>>(bufs[0], maps[0], t0);
    else k_tma<4><<<grid, 256, smem>>>(bufs[0], maps[0], t0);
    CK(cudaDeviceSynchronize());
    std::vector<unsigned long long> tl((size_t)grid * MAXT);
    unsigned long long t0h;
    CK(cudaMemcpy(tl.data(), bufs[0].t_land, tl.size() * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&t0h, t0, 8, cudaMemcpyDeviceToHost));
    printf("n=%d grid=%d issuing threads=%d: %.2f us per launch\n", n, grid, variant == 0 ? 1 : 4, ms * 1e3f / (reps * ring));
    for (int k = 0; k < 4; ++k) {
      std::vector<double> v;
      for (int c = 0; c < grid; ++c)
        if (c + k * grid < tiles) v.push_back((double)(tl[(size_t)c * MAXT + k] - t0h) / 1e3);
      if (v.empty()) continue;
      std::sort(v.begin(), v.end());
      printf("   tile #%d of a CTA (%3zu CTAs): landed median %.2f us  p10 %.2f  p90 %.2f  max %.2f after the first CTA's start\n", k,
             v.size(), v[v.size() / 2], v[v.size() / 10], v[v.size() * 9 / 10], v.back());
    }
  }
  return 0;
}
