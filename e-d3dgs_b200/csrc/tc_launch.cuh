// Host-side launch of the tensor-core GEMM block: picks the instantiation (tile width, prepared
// weights or not), splits weight operands into UMMA images first, orders blocks so that the batch
// entries (heads) of one row tile are adjacent.
#pragma once
#include "tc_gemm.cuh"
#include "tc_rows.cuh"
#include "tc_wgrad.cuh"
#include "fused_mlp.cuh"
#include "tc_wgrad_ts.cuh"
#include "profiler.h"
#include <cstdlib>

namespace ed3dgs {

// floats of scratch the prepared-weight images of one launch can need:
// 5 heads x K=128 or one matrix with K=640, NT <= 128  ->  40 chunks x 128 x 8 float4
inline size_t tc_prep_scratch_floats() { return (size_t)48 * 128 * 8 * 4; }

// Programme of one tile of 256 Gaussians for tc_rows_kernel: the 16-wide K slabs of every batch
// entry in order, with the flags that open and close an accumulation.
inline bool tc_rows_build(TcArgs& a) {
  int ns = 0;
  for (int i = 0; i < a.nbatch; ++i) {
    const int Ki = a.b[i].K > 0 ? a.b[i].K : a.K;
    if (Ki <= 0 || (Ki & 3)) return false;
    for (int k0 = 0; k0 < Ki; k0 += 16) {
      if (ns >= TC_MAX_STEPS) return false;
      const int left = Ki - k0;
      const uint32_t npieces = left >= 16 ? 4 : (left + 3) / 4;
      const uint32_t nk = left > 8 ? 2 : 1;
      const uint32_t flags = (k0 == 0 ? TC_STEP_FIRST : 0) | (k0 + 16 >= Ki ? TC_STEP_LAST : 0);
      a.steps[ns++] = (uint32_t)i | ((uint32_t)(k0 / 16) << 8) | (npieces << 16) | (nk << 20) | (flags << 24);
    }
  }
  a.nsteps = ns;
  a.ntiles = (a.M + ROWS_TG - 1) / ROWS_TG;
  return ns > 0;
}

// SMs the persistent kernels of the deformation network leave free (ed3dgs_deform_set_sm_reserve): a concurrent NCCL
// collective needs a few SMs of its own, otherwise its CTAs wait for a persistent CTA to retire and the CTAs of the
// next persistent launch that lose the race form a second wave (static tile assignment: the kernel takes twice as long).
// Both reserves (and the event in deform.cu) are kept PER DEVICE, keyed by the device that is current at the call: a
// process that drives two GPUs sets and reads them independently.
constexpr int TC_MAX_DEVICES = 64;
inline int tc_device_slot() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= TC_MAX_DEVICES) dev = 0;
  return dev;
}
inline int& tc_sm_reserve() {
  static int reserve[TC_MAX_DEVICES] = {};
  return reserve[tc_device_slot()];
}
// A larger reserve for the weight-gradient launch of the LAST stage only (ed3dgs_deform_set_sm_reserve_last): the exchange
// of the embedding gradient runs next to exactly that launch and has one launch (1.2 ms) to finish in.
inline int& tc_sm_reserve_last() {
  static int reserve[TC_MAX_DEVICES] = {};
  return reserve[tc_device_slot()];
}

inline int tc_sm_count() {   // of the current device (not cached: a process may drive several GPUs)
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
    sms = 148;
  const int r = tc_sm_reserve();
  return (r > 0 && r < sms) ? sms - r : sms;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda).
typedef CUresult (*tc_encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                       const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                       CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline tc_encode_tiled_fn tc_encode_tiled() {
  static tc_encode_tiled_fn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<tc_encode_tiled_fn>(ptr);
  }
  return fn;
}

template <int EPI>
inline cudaError_t tc_rows_run(const TcArgs& a, int grid, cudaStream_t stream) {
  {  // a per-device attribute: set on every launch (a process-wide flag would miss the second GPU)
    const cudaError_t e =
        cudaFuncSetAttribute(tc_rows_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ROWS_SMEM_BYTES);
    if (e != cudaSuccess) return e;
  }
  tc_rows_kernel<EPI><<<grid, ROWS_THREADS, ROWS_SMEM_BYTES, stream>>>(a);
  return cudaGetLastError();
}

inline cudaError_t tc_rows_launch(TcArgs& a, const char* name, cudaStream_t stream) {
  if ((a.lda & 3) || !tc_rows_build(a)) return cudaErrorInvalidValue;
  for (int i = 0; i < a.nbatch; ++i)
    if (a.b[i].n_valid > 128) return cudaErrorInvalidValue;
  // Activation tiles come in by TMA: 256 rows x 16 columns, 64-byte swizzle = the UMMA's K-major SW64
  // layout. Columns >= K_b and rows >= M are zero-filled by the unit.
  tc_encode_tiled_fn enc = tc_encode_tiled();
  if (!enc) return cudaErrorNotSupported;
  for (int i = 0; i < a.nbatch; ++i) {
    const int Ki = a.b[i].K > 0 ? a.b[i].K : a.K;
    if (i > 0 && a.b[i].A == a.b[i - 1].A && Ki == (a.b[i - 1].K > 0 ? a.b[i - 1].K : a.K)) {
      a.xmap[i] = a.xmap[i - 1];
      continue;
    }
    const cuuint64_t dims[2] = {(cuuint64_t)Ki, (cuuint64_t)a.M};
    const cuuint64_t strides[1] = {(cuuint64_t)a.lda * sizeof(float)};
    const cuuint32_t box[2] = {16, (cuuint32_t)ROWS_TG};
    const cuuint32_t estr[2] = {1, 1};
    if ((reinterpret_cast<uintptr_t>(a.b[i].A) & 15) != 0) return cudaErrorInvalidValue;
    const CUresult r = enc(&a.xmap[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(a.b[i].A), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
  }
  const int grid = a.ntiles < tc_sm_count() ? a.ntiles : tc_sm_count();
  ProfScope ps(name, stream);
  switch (a.epilogue) {
    case TC_EPI_BIAS_ACT: return tc_rows_run<TC_EPI_BIAS_ACT>(a, grid, stream);
    case TC_EPI_MASK: return tc_rows_run<TC_EPI_MASK>(a, grid, stream);
    case TC_EPI_ACCUM: return tc_rows_run<TC_EPI_ACCUM>(a, grid, stream);
    case TC_EPI_HEAD: return tc_rows_run<TC_EPI_HEAD>(a, grid, stream);
    default: return cudaErrorInvalidValue;
  }
}

// Weight gradients: D = P^T Q over all Gaussians, scattered into the gradient matrices (tc_wgrad.cuh).
// ts: the kernel of tc_wgrad_ts.cuh (P of every group feature-major and fed through tensor memory, Q at most 128 columns).
inline cudaError_t tc_wgrad_launch(const WgArgs& a_in, cudaStream_t stream, bool ts = false) {
  if (a_in.ngroups <= 0 || a_in.N <= 0) return cudaSuccess;
  if (a_in.ngroups > WG_MAX_GROUPS) return cudaErrorInvalidValue;
  WgArgs a = a_in;
  if (ts)
    for (int i = 0; i < a.ngroups; ++i)
      if (!a.g[i].pk || a.g[i].qcols > 128) return cudaErrorInvalidValue;
  tc_encode_tiled_fn enc = tc_encode_tiled();
  if (!enc) return cudaErrorNotSupported;
  for (int i = 0; i < a.ngroups; ++i) {
    WgGroup& g = a.g[i];
    g.qbox = g.qk ? (g.qcols > 128 ? 256 : (g.qcols + 15) / 16 * 16) : 0;
    if (g.pcols <= 0 || g.pcols > 128 || (g.pcols & 3) || g.qcols <= 0 || g.qcols > 256 || (g.qcols & 3) ||
        (g.ldp & 3) || (g.ldq & 3) || g.nseg < 0 || g.nseg > 2 || (reinterpret_cast<uintptr_t>(g.P) & 15) ||
        (reinterpret_cast<uintptr_t>(g.Q) & 15))
      return cudaErrorInvalidValue;
    // row-major operand: tiles of 32 features x 16 Gaussians, 32-byte chunks swizzled inside 128-byte rows = the UMMA's
    // MN-major layout for 32-bit elements (SWIZZLE_128B_BASE32B); columns >= cols and rows >= N are zero-filled by the unit.
    // feature-major operand ([tile][rows][128 Gaussians]): one box of 16 Gaussians x 128 / 256 feature rows, SWIZZLE_64B
    // = the K-major layout; feature rows >= cols are zero-filled.
    const cuuint32_t estr3[3] = {1, 1, 1};
    const cuuint32_t estr[2] = {1, 1};
    const cuuint64_t ntiles = (cuuint64_t)((a.N + 127) / 128);
    for (int side = 0; side < 2; ++side) {
      const bool km = side ? g.qk != 0 : g.pk != 0;
      const int cols = side ? g.qcols : g.pcols;
      CUtensorMap* map = side ? &a.qmap[i] : &a.pmap[i];
      float* base = const_cast<float*>(side ? g.Q : g.P);
      CUresult r;
      if (km) {
        const long rows_total = side ? g.qrow_total : g.prow_total;
        const cuuint64_t dims[3] = {128, (cuuint64_t)cols, ntiles};
        const cuuint64_t strides[2] = {128 * sizeof(float), (cuuint64_t)rows_total * 128 * sizeof(float)};
        const cuuint32_t box[3] = {16, (cuuint32_t)(side ? g.qbox : 128), 1};
        r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, base, dims, strides, box, estr3, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      } else {
        const cuuint32_t box[2] = {32, 16};
        const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)a.N};
        const cuuint64_t strides[1] = {(cuuint64_t)(side ? g.ldq : g.ldp) * sizeof(float)};
        r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      }
      if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
    }
  }
  {  // a per-device attribute: set on every launch (a process-wide flag would miss the second GPU)
    const cudaError_t e = ts ? cudaFuncSetAttribute(tc_wgrad_ts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WT_SMEM_BYTES)
                             : cudaFuncSetAttribute(tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WG_SMEM_BYTES);
    if (e != cudaSuccess) return e;
  }
  static const int slice_env = [] { const char* e = getenv("ED3DGS_WG_SLICE"); const int v = e ? atoi(e) : 0; return v >= 128 ? v / 128 * 128 : 0; }();
  a.slice = slice_env ? slice_env : WG_SLICE;
  const int items = ((a.N + a.slice - 1) / a.slice) * a.ngroups;
  const int grid = items < tc_sm_count() ? items : tc_sm_count();
  ProfScope ps(ts ? "tc_wgrad_ts_kernel" : "tc_wgrad_kernel", stream);   // the scope carries the name of the kernel that runs
  if (ts) tc_wgrad_ts_kernel<<<grid, WT_THREADS, WT_SMEM_BYTES, stream>>>(a);
  else tc_wgrad_kernel<<<grid, WG_THREADS, WG_SMEM_BYTES, stream>>>(a);
  return cudaGetLastError();
}

inline cudaError_t tc_launch(int amaj, int bmaj, const TcArgs& a_in, int nbatch, float* prep_scratch,
                             cudaStream_t stream) {
  if (nbatch <= 0) return cudaSuccess;
  TcArgs a = a_in;
  a.nbatch = nbatch;
  if (a.M <= 0) return cudaSuccess;
  if (amaj == TC_K_MAJOR) {
    // weights: split once into per-slab hi/lo images of 128 rows (rows >= n_valid are zero)
    a.NT = 128;
    int kmax = 0;
    for (int i = 0; i < nbatch; ++i) kmax = max(kmax, a.b[i].K > 0 ? a.b[i].K : a.K);
    const int chunks = (kmax + TC_KC - 1) / TC_KC;
    const long stride_f4 = (long)chunks * a.NT * (2 * TC_PC);
    if ((size_t)stride_f4 * nbatch * 4 > tc_prep_scratch_floats() || prep_scratch == nullptr) return cudaErrorInvalidValue;
    for (int i = 0; i < nbatch; ++i) a.b[i].Bprep = reinterpret_cast<const float4*>(prep_scratch) + i * stride_f4;
    {
      ProfScope ps("tc_prepare_b", stream);
      const dim3 pg(chunks, nbatch, 1);
      if (bmaj == TC_K_MAJOR) tc_prepare_b<TC_K_MAJOR><<<pg, 256, 0, stream>>>(a, reinterpret_cast<float4*>(prep_scratch), stride_f4);
      else tc_prepare_b<TC_MN_MAJOR><<<pg, 256, 0, stream>>>(a, reinterpret_cast<float4*>(prep_scratch), stride_f4);
    }
    const char* name = (bmaj == TC_MN_MAJOR) ? "tc_rows_kernel<dgrad>" : "tc_rows_kernel<fwd>";
    return tc_rows_launch(a, name, stream);
  }
  return cudaErrorInvalidValue;   // weight gradients go through tc_wgrad_launch
}

}  // namespace ed3dgs
