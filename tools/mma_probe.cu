// Hardware probe for the tcgen05 features the round-2 kernels rely on (run once on a B200, record under profiles/):
//   T1  K-major f16 x f16 (sanity)                      T2  mixed operand types in kind::f16 (A f16, B bf16)
//   T3  MN-major A (bf16)                               T4  MN-major B (bf16)
//   T5  cta_group::1, M = 64: which TMEM lanes/columns hold which (row, column) of D
//   T6  cta_group::2, M = 128 (64 rows per CTA): the same map for both CTAs
//   T7  issue rate of small MMAs (cycles per MMA for N = 16/32/96/192, 1 CTA and CTA pair)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/mma_probe tools/mma_probe.cu
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../3d-pose-estimation_b200/csrc/tc_common.cuh"

namespace smplb {
void set_error(const std::string&) {}
int32_t fail(int32_t code, const std::string&) { return code; }
}  // namespace smplb
using namespace smplb::tc;

#define CK(x)                                                                       \
  do {                                                                              \
    cudaError_t e_ = (x);                                                           \
    if (e_ != cudaSuccess) {                                                        \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(1);                                                                      \
    }                                                                               \
  } while (0)

struct ProbeArgs {
  const uint4* a_img;   // per CTA: a_bytes
  const uint4* b_img;   // per CTA: b_bytes
  int a_bytes, b_bytes;
  uint32_t idesc;
  uint32_t a_lbo, a_sbo, b_lbo, b_sbo;
  uint32_t a_kstep, b_kstep;   // byte advance per K = 16 step
  int ksteps;
  int ncols;      // TMEM columns to dump
  int repeat;     // T7: number of back-to-back MMAs (0: single pass)
  int nacc;       // T7: rotate over this many accumulators (column offset acc_stride each)
  int acc_stride;
  float* out;     // [ctas][128][ncols]
  long long* cycles;
};

__device__ __forceinline__ uint64_t mk_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return umma_desc_kmajor_noswz(addr, lbo, sbo);   // same bit layout for both major modes (no swizzle)
}

__device__ __forceinline__ void umma_f16_ss_pred(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                                 uint32_t accumulate, uint32_t issue) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.ne.b32 q, %5, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(issue)
      : "memory");
}

template <int PAIR>
__global__ void __launch_bounds__(128, 1) probe_kernel(ProbeArgs p) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  uint4* sa = reinterpret_cast<uint4*>(smem);
  uint4* sb = reinterpret_cast<uint4*>(smem + ((p.a_bytes + 127) / 128) * 128);
  const uint32_t rank = PAIR ? cluster_ctarank() : 0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < p.a_bytes / 16; i += blockDim.x) sa[i] = p.a_img[(size_t)rank * (p.a_bytes / 16) + i];
  for (int i = threadIdx.x; i < p.b_bytes / 16; i += blockDim.x) sb[i] = p.b_img[(size_t)rank * (p.b_bytes / 16) + i];
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    if (PAIR) tmem_alloc_pair(&tmem_slot, 512); else tmem_alloc(&tmem_slot, 512);
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  // clear the accumulator area so untouched cells read as a recognisable value
  {
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
    for (int c = 0; c < p.ncols; c += 1) {
      uint32_t v = __float_as_uint(-7777.0f);
      asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr + c), "r"(v) : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  long long t0 = 0, t1 = 0;
  if (warp == 1 && rank == 0) {
    const uint32_t issue = elect_one() ? 1u : 0u;
    const uint32_t ab = smem_u32(sa), bb = smem_u32(sb);
    t0 = clock64();
    const int reps = p.repeat > 0 ? p.repeat : 1;
    if (p.repeat > 0) {
      // tight issue loop: descriptors and accumulator addresses are loop-invariant, 8 MMAs per trip, whole warp converged
      const uint64_t ad = mk_desc(ab, p.a_lbo, p.a_sbo), bd = mk_desc(bb, p.b_lbo, p.b_sbo);
      const uint32_t idesc = p.idesc;
      uint32_t dc[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) dc[i] = tmem + (p.nacc > 1 ? (uint32_t)((i % p.nacc) * p.acc_stride) : 0u);
#pragma unroll 1
      for (int r = 0; r < reps; r += 8) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (PAIR) umma_f16_ss_pair(dc[i], ad, bd, idesc, 1u, issue);
          else umma_f16_ss_pred(dc[i], ad, bd, idesc, 1u, issue);
        }
      }
    } else
    for (int r = 0; r < reps; ++r)
      for (int ks = 0; ks < p.ksteps; ++ks) {
        const uint64_t ad = mk_desc(ab + ks * p.a_kstep, p.a_lbo, p.a_sbo);
        const uint64_t bd = mk_desc(bb + ks * p.b_kstep, p.b_lbo, p.b_sbo);
        const uint32_t accum = (ks != 0) ? 1u : 0u;
        if (PAIR) umma_f16_ss_pair(tmem, ad, bd, p.idesc, accum, issue);
        else if (issue) umma_f16_ss(tmem, ad, bd, p.idesc, accum);
      }
    if (PAIR) umma_commit_pair(smem_u32(&bar), 3, issue);
    else if (issue) umma_commit(&bar);
    __syncwarp();
  }
  if (PAIR) mbar_wait_cluster(&bar, 0); else mbar_wait(&bar, 0);
  if (warp == 1 && rank == 0) {
    t1 = clock64();
    if (lane == 0 && p.cycles) p.cycles[0] = t1 - t0;
  }
  tc_fence_after();
  {
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
    float* o = p.out + ((size_t)rank * 128 + threadIdx.x) * p.ncols;
    for (int c0 = 0; c0 < p.ncols; c0 += 16) {
      uint32_t r[16];
      tmem_ld16(taddr + c0, r);
      tmem_ld_wait();
      for (int j = 0; j < 16 && c0 + j < p.ncols; ++j) o[c0 + j] = __uint_as_float(r[j]);
    }
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_pair(tmem, 512); else tmem_dealloc(tmem, 512);
  }
}

// ---------------- host side ----------------
static uint16_t to16(float x, int bf) {
  if (bf) { __nv_bfloat16 h = __float2bfloat16(x); uint16_t u; memcpy(&u, &h, 2); return u; }
  __half h = __float2half(x); uint16_t u; memcpy(&u, &h, 2); return u;
}
// image of a [rows x K] 16-bit matrix; mn_major = 0: K-major core matrices (8 rows x 16 B), 1: MN-major (8 k x 16 B)
static std::vector<uint16_t> image(const std::vector<float>& m, int rows, int K, int bf, int mn_major, uint32_t lbo,
                                   uint32_t sbo) {
  size_t bytes = 0;
  for (int r = 0; r < rows; ++r)
    for (int k = 0; k < K; ++k) {
      size_t off = (size_t)(r / 8) * sbo + (size_t)(k / 8) * lbo + (mn_major ? (k % 8) * 16 + (r % 8) * 2 : (r % 8) * 16 + (k % 8) * 2);
      if (off + 2 > bytes) bytes = off + 2;
    }
  bytes = (bytes + 127) / 128 * 128;
  std::vector<uint16_t> img(bytes / 2, 0);
  for (int r = 0; r < rows; ++r)
    for (int k = 0; k < K; ++k) {
      size_t off = (size_t)(r / 8) * sbo + (size_t)(k / 8) * lbo + (mn_major ? (k % 8) * 16 + (r % 8) * 2 : (r % 8) * 16 + (k % 8) * 2);
      img[off / 2] = to16(m[(size_t)r * K + k], bf);
    }
  return img;
}
static uint32_t idesc(int afmt, int bfmt, int amaj, int bmaj, int M, int N) {
  return (1u << 4) | ((uint32_t)afmt << 7) | ((uint32_t)bfmt << 10) | ((uint32_t)amaj << 15) | ((uint32_t)bmaj << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

struct Case {
  const char* name;
  int pair, M, N, K, abf, bbf, amn, bmn, repeat, decode, nacc;
};

static void run(const Case& c) {
  const int ctas = c.pair ? 2 : 1;
  const int Mc = c.M / ctas, Nc = c.N / ctas;   // rows of A / B held by each CTA
  std::vector<float> A((size_t)c.M * c.K), B((size_t)c.N * c.K);
  for (int m = 0; m < c.M; ++m)
    for (int k = 0; k < c.K; ++k) A[(size_t)m * c.K + k] = c.decode ? (k == 0 ? (float)m : (k == 1 ? 1.f : 0.f)) : (float)(((m * 3 + k * 5) % 7) - 3);
  for (int n = 0; n < c.N; ++n)
    for (int k = 0; k < c.K; ++k) B[(size_t)n * c.K + k] = c.decode ? (k == 0 ? 1.f : (k == 1 ? 256.f * (n + 1) : 0.f)) : (float)(((n * 2 + k) % 5) - 2);
  // layouts: K-major [k/8][row/8] (LBO = rows*16, SBO = 128); MN-major [row/8][k/8] (LBO = 128, SBO = (K/8)*128)
  const uint32_t a_lbo = c.amn ? 128u : (uint32_t)Mc * 16, a_sbo = c.amn ? (uint32_t)(c.K / 8) * 128 : 128u;
  const uint32_t b_lbo = c.bmn ? 128u : (uint32_t)Nc * 16, b_sbo = c.bmn ? (uint32_t)(c.K / 8) * 128 : 128u;
  std::vector<uint16_t> aimg, bimg;
  int a_bytes = 0, b_bytes = 0;
  for (int r = 0; r < ctas; ++r) {
    std::vector<float> As(A.begin() + (size_t)r * Mc * c.K, A.begin() + (size_t)(r + 1) * Mc * c.K);
    std::vector<float> Bs(B.begin() + (size_t)r * Nc * c.K, B.begin() + (size_t)(r + 1) * Nc * c.K);
    auto ai = image(As, Mc, c.K, c.abf, c.amn, a_lbo, a_sbo), bi = image(Bs, Nc, c.K, c.bbf, c.bmn, b_lbo, b_sbo);
    a_bytes = (int)ai.size() * 2; b_bytes = (int)bi.size() * 2;
    aimg.insert(aimg.end(), ai.begin(), ai.end());
    bimg.insert(bimg.end(), bi.begin(), bi.end());
  }
  ProbeArgs p;
  void *da, *db; float* dout; long long* dcyc;
  const int ncols = c.N < 16 ? 16 : c.N;
  CK(cudaMalloc(&da, aimg.size() * 2)); CK(cudaMalloc(&db, bimg.size() * 2));
  CK(cudaMalloc(&dout, (size_t)ctas * 128 * ncols * 4)); CK(cudaMalloc(&dcyc, 8));
  CK(cudaMemcpy(da, aimg.data(), aimg.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, bimg.data(), bimg.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemset(dcyc, 0, 8));
  p.a_img = (const uint4*)da; p.b_img = (const uint4*)db; p.a_bytes = a_bytes; p.b_bytes = b_bytes;
  p.idesc = idesc(c.abf, c.bbf, c.amn, c.bmn, c.M, c.N);
  p.a_lbo = a_lbo; p.a_sbo = a_sbo; p.b_lbo = b_lbo; p.b_sbo = b_sbo;
  p.a_kstep = c.amn ? 2 * 128u : 2 * a_lbo; p.b_kstep = c.bmn ? 2 * 128u : 2 * b_lbo;
  p.ksteps = c.K / 16; p.ncols = ncols; p.repeat = c.repeat; p.out = dout; p.cycles = dcyc;
  p.nacc = c.nacc; p.acc_stride = c.pair && c.M == 128 ? c.N / 2 : c.N;
  const size_t smem = (size_t)((a_bytes + 127) / 128) * 128 + b_bytes + 256;
  if (c.pair) {
    CK(cudaFuncSetAttribute(probe_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    CK(cudaLaunchKernelEx(&cfg, probe_kernel<1>, p));
  } else {
    CK(cudaFuncSetAttribute(probe_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    probe_kernel<0><<<1, 128, smem>>>(p);
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("%-44s CUDA error: %s\n", c.name, cudaGetErrorString(e)); exit(2); }
  std::vector<float> out((size_t)ctas * 128 * ncols);
  long long cyc = 0;
  CK(cudaMemcpy(out.data(), dout, out.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost));
  if (c.repeat > 0) {
    printf("%-44s %d MMAs (M=%d N=%d K=16%s, %d accumulators): %lld cycles = %.1f per MMA\n", c.name, c.repeat * (c.K / 16), c.M, c.N,
           c.pair ? ", pair" : "", c.nacc > 1 ? c.nacc : 1, cyc, (double)cyc / (c.repeat * (c.K / 16)));
  } else if (c.decode) {
    printf("%s: TMEM map (value = m + 256(n+1)); '.' = untouched\n", c.name);
    for (int r = 0; r < ctas; ++r) {
      // summarise: for every lane, the row m it holds and the set of n columns
      int printed = 0;
      for (int lane = 0; lane < 128; ++lane) {
        int m0 = -1, nmin = 1 << 30, nmax = -1, cnt = 0, cfirst = -1, clast = -1, consistent = 1;
        for (int col = 0; col < ncols; ++col) {
          const float v = out[((size_t)r * 128 + lane) * ncols + col];
          if (v == -7777.0f) continue;
          const int iv = (int)v, m = iv % 256, n = iv / 256 - 1;
          if (m0 < 0) m0 = m; else if (m != m0) consistent = 0;
          if (n < nmin) nmin = n; if (n > nmax) nmax = n;
          if (cfirst < 0) cfirst = col; clast = col; ++cnt;
        }
        if (lane < 4 || lane % 16 == 0 || lane % 16 == 15 || !consistent) {
          if (cnt == 0) printf("  cta %d lane %3d: untouched\n", r, lane);
          else printf("  cta %d lane %3d: row m=%3d%s, n=%d..%d in cols %d..%d (%d cells)\n", r, lane, m0, consistent ? "" : " (MIXED rows)", nmin, nmax, cfirst, clast, cnt);
          ++printed;
        }
      }
    }
  } else {
    double maxerr = 0;
    for (int m = 0; m < c.M; ++m)
      for (int n = 0; n < c.N; ++n) {
        double ref = 0;
        for (int k = 0; k < c.K; ++k) ref += (double)A[(size_t)m * c.K + k] * B[(size_t)n * c.K + k];
        const int r = c.pair ? m / Mc : 0, lane = c.pair ? m % Mc : m;   // pair M=256: lane = row within the CTA
        const double got = out[((size_t)r * 128 + lane) * ncols + n];
        if (fabs(got - ref) > maxerr) maxerr = fabs(got - ref);
      }
    printf("%-44s max |err| = %g  %s\n", c.name, maxerr, maxerr == 0 ? "OK" : "MISMATCH");
  }
  cudaFree(da); cudaFree(db); cudaFree(dout); cudaFree(dcyc);
}

int main() {
  const Case cases[] = {
      {"T1 K-major f16 x f16 M128 N32 K32", 0, 128, 32, 32, 0, 0, 0, 0, 0, 0},
      {"T1b K-major bf16 x bf16 M128 N32 K32", 0, 128, 32, 32, 1, 1, 0, 0, 0, 0},
      // T2 (A f16 x B bf16 in one kind::f16 MMA) was run once: "illegal instruction" -- operand types must match
      {"T3 A MN-major bf16, B K-major bf16", 0, 128, 32, 32, 1, 1, 1, 0, 0, 0},
      {"T4 A K-major bf16, B MN-major bf16", 0, 128, 32, 32, 1, 1, 0, 1, 0, 0},
      {"T4b A MN-major, B MN-major, N=224 K=64", 0, 128, 224, 64, 1, 1, 1, 1, 0, 0},
      {"T4c pair M256 N224: A MN-major, B K-major", 1, 256, 224, 64, 1, 1, 1, 0, 0, 0},
      {"T5 cta_group::1 M=64 N=32 map", 0, 64, 32, 16, 0, 0, 0, 0, 0, 1},
      {"T5b cta_group::1 M=128 N=32 map", 0, 128, 32, 16, 0, 0, 0, 0, 0, 1},
      {"T6 cta_group::2 M=128 N=32 map", 1, 128, 32, 16, 0, 0, 0, 0, 0, 1},
      {"T6b cta_group::2 M=128 N=192 map", 1, 128, 192, 16, 0, 0, 0, 0, 0, 1},
      {"T6c cta_group::2 M=256 N=32 map", 1, 256, 32, 16, 0, 0, 0, 0, 0, 1},
      {"T7 1-CTA M128 N16 1 acc", 0, 128, 16, 16, 0, 0, 0, 0, 512, 0, 1},
      {"T7 1-CTA M128 N16 2 acc", 0, 128, 16, 16, 0, 0, 0, 0, 512, 0, 2},
      {"T7 1-CTA M128 N16 4 acc", 0, 128, 16, 16, 0, 0, 0, 0, 512, 0, 4},
      {"T7 1-CTA M128 N16 8 acc", 0, 128, 16, 16, 0, 0, 0, 0, 512, 0, 8},
      {"T7 1-CTA M128 N32 4 acc", 0, 128, 32, 16, 0, 0, 0, 0, 512, 0, 4},
      {"T7 1-CTA M128 N96 1 acc", 0, 128, 96, 16, 0, 0, 0, 0, 512, 0, 1},
      {"T7 1-CTA M128 N96 2 acc", 0, 128, 96, 16, 0, 0, 0, 0, 512, 0, 2},
      {"T7 1-CTA M128 N96 4 acc", 0, 128, 96, 16, 0, 0, 0, 0, 512, 0, 4},
      {"T7 1-CTA M128 N192 1 acc", 0, 128, 192, 16, 0, 0, 0, 0, 512, 0, 1},
      {"T7 1-CTA M128 N192 2 acc", 0, 128, 192, 16, 0, 0, 0, 0, 512, 0, 2},
      {"T7 1-CTA M128 N256 1 acc", 0, 128, 256, 16, 0, 0, 0, 0, 512, 0, 1},
      {"T7 1-CTA M128 N256 2 acc", 0, 128, 256, 16, 0, 0, 0, 0, 512, 0, 2},
      {"T7 1-CTA M128 N32 K64 (4 ksteps) 1 acc", 0, 128, 32, 64, 0, 0, 0, 0, 128, 0, 1},
      {"T7 1-CTA M128 N32 K64 (4 ksteps) 4 acc", 0, 128, 32, 64, 0, 0, 0, 0, 128, 0, 4},
      {"T7 pair M128 N16 1 acc", 1, 128, 16, 16, 0, 0, 0, 0, 512, 0, 1},
      {"T7 pair M128 N16 4 acc", 1, 128, 16, 16, 0, 0, 0, 0, 512, 0, 4},
      {"T7 pair M128 N32 1 acc", 1, 128, 32, 16, 0, 0, 0, 0, 512, 0, 1},
      {"T7 pair M128 N32 4 acc", 1, 128, 32, 16, 0, 0, 0, 0, 512, 0, 4},
      {"T7 pair M128 N32 8 acc", 1, 128, 32, 16, 0, 0, 0, 0, 512, 0, 8},
      {"T7 pair M128 N192 1 acc", 1, 128, 192, 16, 0, 0, 0, 0, 512, 0, 1},
      {"T7 pair M128 N192 2 acc", 1, 128, 192, 16, 0, 0, 0, 0, 512, 0, 2},
      {"T7 pair M256 N224 1 acc", 1, 256, 224, 16, 0, 0, 0, 0, 512, 0, 1},
      {"T7 pair M256 N224 2 acc", 1, 256, 224, 16, 0, 0, 0, 0, 512, 0, 2},
      {"T7 pair M256 N256 1 acc", 1, 256, 256, 16, 0, 0, 0, 0, 512, 0, 1},
      {"T7 pair M256 N256 2 acc", 1, 256, 256, 16, 0, 0, 0, 0, 512, 0, 2},
  };
  for (const Case& c : cases) run(c);
  return 0;
}
