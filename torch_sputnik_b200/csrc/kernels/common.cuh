// common.cuh -- shared device/host helpers for the tsp kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "tsp kernels are written for sm_100a (B200) only"
#endif

namespace tsp {

constexpr int kWarp = 32;
constexpr unsigned kFullMask = 0xffffffffu;

// ---------------------------------------------------------------- host side
void set_error(const char* fmt, ...);  // defined in tsp_abi.cu (thread-local message)
extern std::atomic<unsigned long long> g_launches;

inline int cuda_fail(cudaError_t e, const char* what) {
  set_error("%s: %s", what, cudaGetErrorString(e));
  return 2;  // TSP_ERR_CUDA
}

#define TSP_CHECK_LAUNCH(what)                                  \
  do {                                                          \
    ::tsp::g_launches.fetch_add(1, std::memory_order_relaxed);  \
    cudaError_t e__ = cudaGetLastError();                       \
    if (e__ != cudaSuccess) return ::tsp::cuda_fail(e__, what); \
  } while (0)

inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int sm_count();  // cached cudaDevAttrMultiProcessorCount of the current device

// -------------------------------------------------------------- device side
#ifdef __CUDACC__

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// Read-only (non-coherent) loads: operands are never written by the same kernel.
__device__ __forceinline__ float ldg_f(const float* p) { return __ldg(p); }
__device__ __forceinline__ int ldg_i(const int* p) { return __ldg(p); }
__device__ __forceinline__ float4 ldg_f4(const float* p) {
  return __ldg(reinterpret_cast<const float4*>(p));
}

// Streaming loads/stores for data touched exactly once (keeps L1/L2 for the gathered operand).
__device__ __forceinline__ float ld_stream_f(const float* p) {
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float4 ld_stream_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ uint4 ld_stream_u4(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream_u4(void* p, uint4 v) {
  asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void st_stream_f(float* p, float v) {
  asm volatile("st.global.cs.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void st_stream_f4(float* p, float4 v) {
  asm volatile("st.global.cs.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

// Packed 2 x fp32 FMA (SASS FFMA2 on sm_100a): d = a * b + c lane-wise on register pairs.
// Same per-lane IEEE rounding as FFMA; halves the issue slots of the FMA stream.
__device__ __forceinline__ void ffma2(float& d0, float& d1, float a0, float a1, float b0,
                                      float b1) {
  unsigned long long a, b, c;
  asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(b0), "f"(b1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(c) : "f"(d0), "f"(d1));
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(c) : "l"(a), "l"(b));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(c));
}

// Butterflies run 1,2,4,8,16 -- the order of sparse_softmax.cu.cc:46-49,57-59 -- so sums round
// the same way as the reference's.
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) v = fmaxf(v, __shfl_xor_sync(kFullMask, v, s));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) v += __shfl_xor_sync(kFullMask, v, s);
  return v;
}

// cp.async (LDGSTS) 16-byte copy global -> shared, L2-only (.cg) so the staged tile
// does not pollute L1.
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
  asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}


// ---- mbarrier + bulk async copy (TMA engine, descriptor-free; SASS: UBLKCP / SYNCS) ----------
__device__ __forceinline__ unsigned smem_u32(const void* p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// Non-blocking phase test: true once the phase with the given parity has completed.
__device__ __forceinline__ bool mbar_test(unsigned long long* bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// global -> shared bulk copy of `bytes` (multiple of 16, both addresses 16-byte aligned),
// completion reported to `bar` as transaction bytes.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, unsigned bytes,
                                         unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// TMA tiled load of a 3-D box (coordinates innermost first) described by a CUtensorMap that
// lives in kernel parameter space (__grid_constant__); completion -> mbarrier transaction bytes.
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const void* tmap, int c0, int c1, int c2,
                                            unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(smem_dst)),
      "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
      : "memory");
}

// L2-aware rasterisation of a 2-D grid of tiles (rows x cols) walked as a 1-D sequence: tiles are
// numbered band by band (srb tile rows per band), inside a band group by group (scb tile columns
// per group), inside a group row-major.  With srb * scb ~ the number of CTAs resident at once, the
// CTAs in flight share ONE group: srb tile rows of the left operand + scb tile columns of the right
// operand -- a footprint sized to stay in L2 -- and successive groups of a band reuse its rows.
__device__ __forceinline__ void raster_tile(int t, int rows, int cols, int srb, int scb, int& r, int& c) {
  const int band_full = srb * cols;
  const int b = t / band_full, tb = t - b * band_full;
  const int band_rows = min(srb, rows - b * srb);
  const int group_full = band_rows * scb;
  const int g = tb / group_full, tg = tb - g * group_full;
  const int w = min(scb, cols - g * scb);
  const int rl = tg / w;
  r = b * srb + rl;
  c = g * scb + (tg - rl * w);
}

// ---- many masks in one launch (the *_many_mask ops: tests/test_attention_many_masks.py:120-150,
// tests/transformer/functions.py of the reference) -------------------------------------------------
// b CSR masks of one shape share a launch: replica r uses mask r / heads.  Mask i's topology is
// row_offsets + i * (m + 1) (LOCAL offsets), column_indices + prefix[i], row_indices + i * m;
// values / sampled outputs are ragged by replica: replica r's nnz_i values start at
// heads * prefix[i] + (r % heads) * nnz_i.  prefix == null: one shared topology (the usual case).
struct MaskTable {
  const int* prefix = nullptr;   // device int[b + 1]: exclusive prefix of the masks' nonzero counts
  int heads = 1;
};
struct MaskView {
  long long topo_rows;    // offset into row_offsets / row_indices, in rows (mask * (m + 1) resp. mask * m)
  long long cols;         // offset into column_indices
  long long values;       // offset of this replica's values / sampled outputs
  int nnz;                // nonzeros of this replica's mask
  int mask;
};
__device__ __forceinline__ MaskView mask_view(const MaskTable& t, int rep, int nnz_shared) {
  MaskView v;
  if (t.prefix == nullptr) {
    v.mask = 0; v.topo_rows = 0; v.cols = 0; v.nnz = nnz_shared; v.values = (long long)rep * nnz_shared;
  } else {
    v.mask = rep / t.heads;
    const int p0 = __ldg(t.prefix + v.mask);
    v.nnz = __ldg(t.prefix + v.mask + 1) - p0;
    v.cols = p0;
    v.values = (long long)t.heads * p0 + (long long)(rep - v.mask * t.heads) * v.nnz;
    v.topo_rows = v.mask;   // scaled by (m + 1) or m at the point of use
  }
  return v;
}

// ---- storage types: fp32, or fp16 / bf16 storage with fp32 arithmetic (the
// reference's half SpMM, cuda_spmm.cu.cc:511-570, keeps fp32 accumulators the same way) -------------
template <typename T>
struct Elem;
template <>
struct Elem<float> {
  static constexpr int kPer16 = 4;   // elements per 16-byte vector
  __device__ __forceinline__ static float to_f(float x) { return x; }
  __device__ __forceinline__ static float from_f(float x) { return x; }
};
template <>
struct Elem<__half> {
  static constexpr int kPer16 = 8;
  __device__ __forceinline__ static float to_f(__half x) { return __half2float(x); }
  __device__ __forceinline__ static __half from_f(float x) { return __float2half_rn(x); }
};
template <>
struct Elem<__nv_bfloat16> {
  static constexpr int kPer16 = 8;
  __device__ __forceinline__ static float to_f(__nv_bfloat16 x) { return __bfloat162float(x); }
  __device__ __forceinline__ static __nv_bfloat16 from_f(float x) { return __float2bfloat16_rn(x); }
};

// two packed 16-bit elements -> two floats
template <typename T>
__device__ __forceinline__ float2 unpack_pair(unsigned w);
template <>
__device__ __forceinline__ float2 unpack_pair<__half>(unsigned w) {
  return __half22float2(*reinterpret_cast<const __half2*>(&w));
}
template <>
__device__ __forceinline__ float2 unpack_pair<__nv_bfloat16>(unsigned w) {
  return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));   // bf16 = top half of fp32
}
template <typename T>
__device__ __forceinline__ unsigned pack_pair(float lo, float hi);
template <>
__device__ __forceinline__ unsigned pack_pair<__half>(float lo, float hi) {
  const __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<const unsigned*>(&h);
}
template <>
__device__ __forceinline__ unsigned pack_pair<__nv_bfloat16>(float lo, float hi) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const unsigned*>(&h);
}

// Mixed-precision FMA (PTX fma.rn.f32.f16 / .bf16, SASS FHFMA[.BF16] on sm_100a): both factors are 16-bit
// elements read straight out of a packed register half, the addend and the result are fp32, ONE
// rounding -- bit-identical to widening both factors and an fp32 FMA, without the two conversion
// instructions per element that made the 16-bit kernels issue-bound (profiles/r02_c4_sddmm_rows_f16_*).
// acc += a.lo * b.lo + a.hi * b.hi over one packed pair, in that order.
template <typename T>
__device__ __forceinline__ float fma_pair_16(unsigned a, unsigned b, float acc);
template <>
__device__ __forceinline__ float fma_pair_16<__half>(unsigned a, unsigned b, float acc) {
  asm("{\n.reg .b16 al, ah, bl, bh;\nmov.b32 {al, ah}, %1;\nmov.b32 {bl, bh}, %2;\n"
      "fma.rn.f32.f16 %0, al, bl, %0;\nfma.rn.f32.f16 %0, ah, bh, %0;\n}"
      : "+f"(acc) : "r"(a), "r"(b));
  return acc;
}
template <>
__device__ __forceinline__ float fma_pair_16<__nv_bfloat16>(unsigned a, unsigned b, float acc) {
  asm("{\n.reg .b16 al, ah, bl, bh;\nmov.b32 {al, ah}, %1;\nmov.b32 {bl, bh}, %2;\n"
      "fma.rn.f32.bf16 %0, al, bl, %0;\nfma.rn.f32.bf16 %0, ah, bh, %0;\n}"
      : "+f"(acc) : "r"(a), "r"(b));
  return acc;
}
// (acc0, acc1) += w * (x.lo, x.hi): one 16-bit factor `w` (low half of the register) against a packed pair.
template <typename T>
__device__ __forceinline__ void fma_scalar_pair_16(unsigned w, unsigned x, float& acc0, float& acc1);
template <>
__device__ __forceinline__ void fma_scalar_pair_16<__half>(unsigned w, unsigned x, float& acc0, float& acc1) {
  asm("{\n.reg .b16 wl, wh, xl, xh;\nmov.b32 {wl, wh}, %2;\nmov.b32 {xl, xh}, %3;\n"
      "fma.rn.f32.f16 %0, wl, xl, %0;\nfma.rn.f32.f16 %1, wl, xh, %1;\n}"
      : "+f"(acc0), "+f"(acc1) : "r"(w), "r"(x));
}
template <>
__device__ __forceinline__ void fma_scalar_pair_16<__nv_bfloat16>(unsigned w, unsigned x, float& acc0, float& acc1) {
  asm("{\n.reg .b16 wl, wh, xl, xh;\nmov.b32 {wl, wh}, %2;\nmov.b32 {xl, xh}, %3;\n"
      "fma.rn.f32.bf16 %0, wl, xl, %0;\nfma.rn.f32.bf16 %1, wl, xh, %1;\n}"
      : "+f"(acc0), "+f"(acc1) : "r"(w), "r"(x));
}

// Raw register image of a vector load of BYTES bytes.
template <int BYTES>
struct RawVec;
template <>
struct RawVec<16> { using type = uint4; };
template <>
struct RawVec<8> { using type = uint2; };
template <>
struct RawVec<4> { using type = unsigned; };
template <>
struct RawVec<2> { using type = unsigned short; };

// E consecutive elements of T at p (aligned to E * sizeof(T)): one read-only vector load, kept as
// raw bits so that several loads can be in flight without holding their fp32 expansion.
template <typename T, int E>
__device__ __forceinline__ typename RawVec<E * sizeof(T)>::type ldg_raw(const T* p) {
  return __ldg(reinterpret_cast<const typename RawVec<E * sizeof(T)>::type*>(p));
}

template <typename T, int E>
__device__ __forceinline__ void unpack_raw(float (&dst)[E], const typename RawVec<E * sizeof(T)>::type& v) {
  if constexpr (sizeof(T) == 4) {
    static_assert(E == 1 || E == 2 || E == 4, "fp32: 1, 2 or 4 elements per load");
    if constexpr (E == 4) {
      dst[0] = __uint_as_float(v.x); dst[1] = __uint_as_float(v.y); dst[2] = __uint_as_float(v.z); dst[3] = __uint_as_float(v.w);
    } else if constexpr (E == 2) {
      dst[0] = __uint_as_float(v.x); dst[1] = __uint_as_float(v.y);
    } else {
      dst[0] = __uint_as_float(v);
    }
  } else {
    static_assert(E == 1 || E == 2 || E == 4 || E == 8, "16-bit: 1, 2, 4 or 8 elements per load");
    if constexpr (E == 8) {
      const float2 a = unpack_pair<T>(v.x), b = unpack_pair<T>(v.y), c = unpack_pair<T>(v.z), d = unpack_pair<T>(v.w);
      dst[0] = a.x; dst[1] = a.y; dst[2] = b.x; dst[3] = b.y; dst[4] = c.x; dst[5] = c.y; dst[6] = d.x; dst[7] = d.y;
    } else if constexpr (E == 4) {
      const float2 a = unpack_pair<T>(v.x), b = unpack_pair<T>(v.y);
      dst[0] = a.x; dst[1] = a.y; dst[2] = b.x; dst[3] = b.y;
    } else if constexpr (E == 2) {
      const float2 a = unpack_pair<T>(v);
      dst[0] = a.x; dst[1] = a.y;
    } else {
      dst[0] = unpack_pair<T>((unsigned)v).x;
    }
  }
}

template <typename T, int E>
__device__ __forceinline__ void load_elems(float (&dst)[E], const T* p) {
  unpack_raw<T, E>(dst, ldg_raw<T, E>(p));
}

// fp32 -> E consecutive elements of T at p (round to nearest even), one vector store.
template <typename T, int E>
__device__ __forceinline__ void store_elems(T* p, const float (&src)[E]) {
  if constexpr (sizeof(T) == 4) {
    if constexpr (E == 4) *reinterpret_cast<float4*>(p) = make_float4(src[0], src[1], src[2], src[3]);
    else if constexpr (E == 2) *reinterpret_cast<float2*>(p) = make_float2(src[0], src[1]);
    else *reinterpret_cast<float*>(p) = src[0];
  } else {
    if constexpr (E == 8)
      *reinterpret_cast<uint4*>(p) = make_uint4(pack_pair<T>(src[0], src[1]), pack_pair<T>(src[2], src[3]),
                                                pack_pair<T>(src[4], src[5]), pack_pair<T>(src[6], src[7]));
    else if constexpr (E == 4)
      *reinterpret_cast<uint2*>(p) = make_uint2(pack_pair<T>(src[0], src[1]), pack_pair<T>(src[2], src[3]));
    else if constexpr (E == 2)
      *reinterpret_cast<unsigned*>(p) = pack_pair<T>(src[0], src[1]);
    else
      *p = Elem<T>::from_f(src[0]);
  }
}

#endif  // __CUDACC__

}  // namespace tsp
