// plan.cuh -- tsp_plan: everything about ONE static CSR topology that does not depend on the
// operand values, computed once instead of on every operator call.
//
// The reference fixes its topology at module construction (modules/sparse_attention.py:49-52,
// modules/sparse_linear.py:78-83) and then calls spmm / sddmm / csr_transpose on it every step;
// its binding redoes all topology work per call (a cuSPARSE buffer query + csr2csc per backward,
// src/transpose_cuda.cu:10-43,90-99).  A plan owns, per topology:
//   * the sortedness verdict the staged kernels need (ascending columns inside every row): one
//     check kernel at creation, the device word copied to pinned host memory behind an event, so
//     later calls know it ON THE HOST -- no per-call check launch, no guarded fallback launch;
//   * the block-start views of the sampled SDDMM (one per tile width) and the occupancy-bitmap
//     view of the tiled SDDMM, built on first use;
//   * the transposed topology + permutation (a child plan), built on first use.
// Unplanned C-ABI calls (tsp_spmm_f32, tsp_sddmm_f32, ...) run on an EPHEMERAL plan on the
// caller's stack: same code path, views in stream-ordered scratch released at the end of the
// call, sortedness unknown to the host (device-guarded staged kernel + guarded gather fallback).
#pragma once
#include <mutex>
#include <vector>

#include "csr_transpose.cuh"
#include "sddmm_sampled.cuh"
#include "spmm_staged.cuh"

struct tsp_plan {
  int m = 0, n = 0, nnz = 0;
  const int* row_offsets = nullptr;      // borrowed: must outlive the plan
  const int* column_indices = nullptr;   // borrowed
  bool ephemeral = false;

  unsigned* d_flag = nullptr;        // device word: 1 = some row's columns are not ascending
  unsigned* h_flag = nullptr;        // pinned host copy (persistent plans)
  cudaEvent_t ev_flag = nullptr;     // recorded behind the device -> host copy
  cudaStream_t flag_stream = nullptr;
  bool check_pending = false;        // ephemeral: flag allocated and zeroed, check not launched yet
  std::atomic<int> sorted{-1};       // host verdict: -1 unknown, 0 unsorted, 1 sorted

  struct View {
    int key;                // > 0: block starts for tile width `key`; kBitmapKey: 64-column bitmaps
    void* ptr;
    cudaEvent_t ev;         // recorded behind the build (persistent plans)
    cudaStream_t stream;    // stream the build ran on
    bool settled;           // build known complete: no cross-stream ordering needed any more
  };
  std::vector<View> views;
  std::mutex mu;

  // transposed topology (persistent plans; built by plan_transposed)
  tsp_plan* transposed = nullptr;
  int* ro_t = nullptr;
  int* ci_t = nullptr;
  int* perm = nullptr;               // values_t[p] = values[perm[p]]
  int* pv_src = nullptr;             // tiled view of perm (csr_transpose.cuh): sources of every output tile in
  unsigned short* pv_loc = nullptr;  // ascending order + their slot inside the tile; null for small topologies
  cudaEvent_t ev_t = nullptr;
  cudaStream_t t_stream = nullptr;
  bool t_settled = false;
};

namespace tsp {

constexpr int kBitmapKey = -64;

inline int check_grid(int m) { return min(ceil_div(m, 8), 4 * sm_count()); }

// Persistent plan: allocate the flag, run the check, start the copy of the verdict to the host.
inline int plan_init_persistent(tsp_plan* p, cudaStream_t stream) {
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&p->d_flag), 256);
  if (e != cudaSuccess) return cuda_fail(e, "tsp_plan_create (flag)");
  e = cudaMallocHost(reinterpret_cast<void**>(&p->h_flag), sizeof(unsigned));
  if (e != cudaSuccess) return cuda_fail(e, "tsp_plan_create (pinned flag)");
  e = cudaEventCreateWithFlags(&p->ev_flag, cudaEventDisableTiming);
  if (e != cudaSuccess) return cuda_fail(e, "tsp_plan_create (event)");
  *p->h_flag = 0xffffffffu;
  p->flag_stream = stream;
  cudaMemsetAsync(p->d_flag, 0, sizeof(unsigned), stream);
  if (p->m > 0 && p->nnz > 0) {
    csr_check_sorted_kernel<<<check_grid(p->m), 256, 0, stream>>>(p->row_offsets, p->column_indices, p->m,
                                                                  p->d_flag);
    TSP_CHECK_LAUNCH("tsp_plan_create (sorted check)");
  }
  cudaMemcpyAsync(p->h_flag, p->d_flag, sizeof(unsigned), cudaMemcpyDeviceToHost, stream);
  e = cudaEventRecord(p->ev_flag, stream);
  if (e != cudaSuccess) return cuda_fail(e, "tsp_plan_create (event record)");
  return 0;
}

inline bool stream_is_capturing(cudaStream_t stream) {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &st) != cudaSuccess) {
    (void)cudaGetLastError();
    return false;
  }
  return st != cudaStreamCaptureStatusNone;
}

// 1 sorted / 0 unsorted / -1 not known on the host (yet).  `stream`: the stream of the call that asks;
// while it is being captured into a CUDA graph the event is NOT queried (cudaEventQuery is one of the
// calls that invalidate a capture) -- the verdict then simply stays unknown for this call.
inline int plan_sorted_state(tsp_plan* p, cudaStream_t stream = nullptr, bool have_stream = false) {
  int s = p->sorted.load(std::memory_order_acquire);
  if (s >= 0 || p->ephemeral || p->ev_flag == nullptr) return s;
  if (have_stream && stream_is_capturing(stream)) return -1;
  if (cudaEventQuery(p->ev_flag) != cudaSuccess) {
    (void)cudaGetLastError();   // cudaErrorNotReady is not sticky, but keep the error state clean
    return -1;
  }
  s = *reinterpret_cast<volatile unsigned*>(p->h_flag) == 0u ? 1 : 0;
  p->sorted.store(s, std::memory_order_release);
  return s;
}

enum { kGuardSorted = 0, kGuardUnknown = 1, kGuardUnsorted = 2 };

// Guard for a staged launch on `stream`.
//   kGuardSorted:   columns known sorted -> guard.flag = null, no fallback launch.
//   kGuardUnknown:  guard.flag = the plan's device word; queue the guarded gather launch behind.
//   kGuardUnsorted: do not launch a staged kernel at all.
inline int plan_guard(tsp_plan* p, cudaStream_t stream, StagedGuard* g, int* rc) {
  *rc = 0;
  g->flag = nullptr;
  const int s = plan_sorted_state(p, stream, true);
  if (s == 1) return kGuardSorted;
  if (s == 0) return kGuardUnsorted;
  if (p->ephemeral) {
    if (p->d_flag == nullptr) {
      cudaError_t e = cudaMallocAsync(reinterpret_cast<void**>(&p->d_flag), 256, stream);
      if (e != cudaSuccess) {
        *rc = cuda_fail(e, "tsp (sortedness flag scratch)");
        return kGuardUnsorted;
      }
      cudaMemsetAsync(p->d_flag, 0, sizeof(unsigned), stream);
      p->check_pending = true;
    }
  } else if (stream != p->flag_stream) {
    // The check ran on another stream.  A stream under capture must not wait on an event recorded
    // outside the capture (it would invalidate the graph): take the kernels that need no verdict.
    if (stream_is_capturing(stream)) return kGuardUnsorted;
    cudaStreamWaitEvent(stream, p->ev_flag, 0);
  }
  g->flag = p->d_flag;
  return kGuardUnknown;
}

// Ephemeral plans only: run the stand-alone check if no view build has carried it.
inline int plan_flush_check(tsp_plan* p, cudaStream_t stream) {
  if (!p->check_pending) return 0;
  p->check_pending = false;
  csr_check_sorted_kernel<<<check_grid(p->m), 256, 0, stream>>>(p->row_offsets, p->column_indices, p->m,
                                                                p->d_flag);
  TSP_CHECK_LAUNCH("tsp (sorted check)");
  return 0;
}

// The view `key` of the plan's topology, built on first use.  Returns null with *rc == 0 when the
// view cannot be built right now (a persistent plan asked for a new view while its stream is being
// captured into a CUDA graph: cudaMalloc is illegal there) -- the caller then uses a kernel that
// needs no view; null with *rc != 0 on error.
inline const void* plan_view(tsp_plan* p, int key, cudaStream_t stream, int* rc) {
  *rc = 0;
  std::lock_guard<std::mutex> lock(p->mu);
  for (auto& v : p->views) {
    if (v.key != key) continue;
    if (!v.settled) {
      if (v.stream == stream) {
        // same stream: ordered by construction
      } else if (stream_is_capturing(stream)) {
        return nullptr;   // neither an event query nor a wait on an outside event is legal under capture
      } else if (cudaEventQuery(v.ev) == cudaSuccess) {
        v.settled = true;
      } else {
        (void)cudaGetLastError();
        cudaStreamWaitEvent(stream, v.ev, 0);
      }
    }
    return v.ptr;
  }
  const size_t bytes = key == kBitmapKey ? sddmm_tiled_view_bytes(p->m, ceil_div(p->n, 64))
                                         : sddmm_sampled_view_bytes(p->m, p->n, key);
  void* ws = nullptr;
  cudaError_t e;
  if (p->ephemeral) {
    e = cudaMallocAsync(&ws, bytes, stream);
  } else {
    if (stream_is_capturing(stream)) return nullptr;
    e = cudaMalloc(&ws, bytes);
  }
  if (e != cudaSuccess) {
    *rc = cuda_fail(e, "tsp (topology view)");
    return nullptr;
  }
  unsigned* flag = p->check_pending ? p->d_flag : nullptr;   // ephemeral: fuse the check into the build
  p->check_pending = false;
  const int brc = key == kBitmapKey
                      ? build_sddmm_tiled_view(p->row_offsets, p->column_indices, p->m, ceil_div(p->n, 64), ws, flag, stream)
                      : build_sddmm_sampled_view(p->row_offsets, p->column_indices, p->m, p->n, key, ws, flag, stream);
  if (brc != 0) {   // the build did not launch: do not keep a view nobody filled
    if (p->ephemeral) cudaFreeAsync(ws, stream); else cudaFree(ws);
    *rc = brc;
    return nullptr;
  }
  tsp_plan::View v{key, ws, nullptr, stream, p->ephemeral};
  if (!p->ephemeral) {
    if (cudaEventCreateWithFlags(&v.ev, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventRecord(v.ev, stream) != cudaSuccess) {
      (void)cudaGetLastError();
      cudaStreamSynchronize(stream);   // no event: make the build visible the blunt way
      v.settled = true;
    }
  }
  p->views.push_back(v);
  *rc = brc;
  return brc == 0 ? ws : nullptr;
}

// Ephemeral plan: hand the stream-ordered scratch back (the kernels using it are already queued).
inline void plan_release_ephemeral(tsp_plan* p, cudaStream_t stream) {
  for (auto& v : p->views) cudaFreeAsync(v.ptr, stream);
  p->views.clear();
  if (p->d_flag) cudaFreeAsync(p->d_flag, stream);
  p->d_flag = nullptr;
}

inline void plan_destroy(tsp_plan* p) {
  if (!p) return;
  if (p->transposed) plan_destroy(p->transposed);
  for (auto& v : p->views) {
    cudaFree(v.ptr);
    if (v.ev) cudaEventDestroy(v.ev);
  }
  if (p->ro_t) cudaFree(p->ro_t);
  if (p->ci_t) cudaFree(p->ci_t);
  if (p->perm) cudaFree(p->perm);
  if (p->pv_src) cudaFree(p->pv_src);
  if (p->pv_loc) cudaFree(p->pv_loc);
  if (p->ev_t) cudaEventDestroy(p->ev_t);
  if (p->d_flag) cudaFree(p->d_flag);
  if (p->h_flag) cudaFreeHost(p->h_flag);
  if (p->ev_flag) cudaEventDestroy(p->ev_flag);
  (void)cudaGetLastError();
  delete p;
}

}  // namespace tsp
