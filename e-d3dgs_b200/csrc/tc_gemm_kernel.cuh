// The GEMM kernel itself (included by tc_gemm.cuh; helpers and the pipeline description live there).
//   NT_T  : compile-time tile width (0 = runtime p.NT)
//   BPREP : B comes as prepared per-chunk UMMA images (weights; tc_prepare_b) and is copied
//           straight into a 3-slot ring with cp.async; only A goes through raw ring + conversion.
#pragma once

namespace ed3dgs {

template <int AMAJ>
__device__ __forceinline__ void tc_epilogue_row(const TcArgs& p, const TcBatch& bt, float* v, int n0, int row, long grow) {
  if (AMAJ == TC_K_MAJOR) {
    const bool full = (n0 + 16 <= bt.n_valid);
    switch (p.epilogue) {
      case TC_EPI_BIAS_ACT: {
        if (full && bt.bias != nullptr) {
          const float4* b4 = reinterpret_cast<const float4*>(bt.bias + n0);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 b = __ldg(b4 + i);
            v[4 * i] += b.x; v[4 * i + 1] += b.y; v[4 * i + 2] += b.z; v[4 * i + 3] += b.w;
          }
        } else if (bt.bias != nullptr) {
          for (int i = 0; i < 16; ++i)
            if (n0 + i < bt.n_valid) v[i] += bt.bias[n0 + i];
        }
        if (p.relu) {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
        }
        if (full && (bt.ldc & 3) == 0) {
          float4* dst = reinterpret_cast<float4*>(bt.C + grow * bt.ldc + n0);
#pragma unroll
          for (int i = 0; i < 4; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        } else {
          for (int i = 0; i < 16; ++i)
            if (n0 + i < bt.n_valid) bt.C[grow * bt.ldc + n0 + i] = v[i];
        }
      } break;
      case TC_EPI_MASK: {
        if (full && (bt.ldc & 3) == 0 && (bt.ldaux & 3) == 0) {
          const float4* msk = reinterpret_cast<const float4*>(bt.aux + grow * bt.ldaux + n0);
          float4* dst = reinterpret_cast<float4*>(bt.C + grow * bt.ldc + n0);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 m = msk[i];
            dst[i] = make_float4(m.x > 0.f ? v[4 * i] : 0.f, m.y > 0.f ? v[4 * i + 1] : 0.f,
                                 m.z > 0.f ? v[4 * i + 2] : 0.f, m.w > 0.f ? v[4 * i + 3] : 0.f);
          }
        } else {
          for (int i = 0; i < 16; ++i) {
            const int n = n0 + i;
            if (n < bt.n_valid) bt.C[grow * bt.ldc + n] = (bt.aux[grow * bt.ldaux + n] > 0.f) ? v[i] : 0.f;
          }
        }
      } break;
      case TC_EPI_ACCUM: {
        if (full && (bt.ldc & 3) == 0) {
          float4* dst = reinterpret_cast<float4*>(bt.C + grow * bt.ldc + n0);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float4 o = p.accumulate ? dst[i] : make_float4(0.f, 0.f, 0.f, 0.f);
            o.x += v[4 * i]; o.y += v[4 * i + 1]; o.z += v[4 * i + 2]; o.w += v[4 * i + 3];
            dst[i] = o;
          }
        } else {
          for (int i = 0; i < 16; ++i) {
            const int n = n0 + i;
            if (n < bt.n_valid) {
              float* dst = bt.C + grow * bt.ldc + n;
              *dst = (p.accumulate ? *dst : 0.f) + v[i];
            }
          }
        }
      } break;
      case TC_EPI_HEAD: {
        for (int i = 0; i < 16; ++i) {
          const int n = n0 + i;
          if (n < bt.n_valid) bt.C[grow * bt.ldc + n] = bt.aux[grow * bt.ldaux + n] + (v[i] + bt.bias[n]) * bt.scale;
        }
      } break;
      default: break;
    }
  } else {
    // weight gradients: D row = feature m, contraction was over Gaussians
    for (int i = 0; i < 16; ++i) {
      const int n = n0 + i;
      if (n < bt.n_valid) {
        float* dst = (p.epilogue == TC_EPI_ATOMIC_T) ? (bt.C + (long)n * bt.ldc + row) : (bt.C + (long)row * bt.ldc + n);
        atomicAdd(dst, v[i]);
      } else if (n == p.ones_col && bt.bias_grad != nullptr) {
        atomicAdd(bt.bias_grad + row, v[i]);
      }
    }
  }
}

template <int AMAJ, int BMAJ, int NT_T, bool BPREP>
__global__ void __launch_bounds__(TC_THREADS, (TC_KC == 8 && BPREP) ? 4 : 2) tc_gemm_kernel(const TcArgs p) {
  extern __shared__ __align__(128) unsigned char tc_smem[];
  __shared__ __align__(8) uint64_t s_bar[2];
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tile = blockIdx.x / p.nbatch;
  const TcBatch bt = p.b[blockIdx.x - tile * p.nbatch];
  const int NT = NT_T ? NT_T : p.NT;
  const int RS = BPREP ? 2 : (NT <= 128 ? 3 : 2);          // raw ring depth
  const int Ktot = bt.K > 0 ? bt.K : p.K;
  const int Kb = bt.kb_valid > 0 ? min(bt.kb_valid, Ktot) : Ktot;

  // shared-memory carve-up
  const int a_raw = TC_M * TC_KC, b_raw = BPREP ? 0 : NT * TC_KC;            // floats per raw slot
  const uint32_t a_f4 = TC_M * (TC_KC / 4), b_f4 = (uint32_t)NT * (TC_KC / 4);
  float* s_raw = reinterpret_cast<float*>(tc_smem);
  float4* s_umma = reinterpret_cast<float4*>(s_raw + (size_t)RS * (a_raw + b_raw));
  const uint32_t stage_f4 = BPREP ? 2 * a_f4 : 2 * (a_f4 + b_f4);
  float4* s_bring = s_umma + 2 * (size_t)stage_f4;                           // BPREP: 3 slots of NT*8 float4
  auto rawA = [&](int slot) { return s_raw + (size_t)slot * (a_raw + b_raw); };
  auto rawB = [&](int slot) { return s_raw + (size_t)slot * (a_raw + b_raw) + a_raw; };
  auto sAhi = [&](int s) { return s_umma + (size_t)s * stage_f4; };
  auto sAlo = [&](int s) { return s_umma + (size_t)s * stage_f4 + a_f4; };
  auto sBhi = [&](int s, int c) { return BPREP ? s_bring + (size_t)(c % 3) * (2 * b_f4) : s_umma + (size_t)s * stage_f4 + 2 * a_f4; };
  auto sBlo = [&](int s, int c) { return sBhi(s, c) + b_f4; };

  uint32_t ncols = 32;
  while (ncols < (uint32_t)NT) ncols <<= 1;
  if (warp == 0) tmem_alloc(&s_tmem, ncols);
  if (tid == 0) {
    mbar_init(&s_bar[0], 1);
    mbar_init(&s_bar[1], 1);
    fence_mbar_init();
  }

  // Tile origin and contraction range.
  int m0 = 0, kbeg = 0, kend = Ktot;
  const float* Abase = bt.A;
  if (AMAJ == TC_K_MAJOR) {
    m0 = tile * TC_M;
    Abase += (long)m0 * p.lda;
  } else {
    kbeg = tile * p.k_per_cta;
    kend = min(Ktot, kbeg + p.k_per_cta);
  }
  const int a_rows_valid = (AMAJ == TC_K_MAJOR) ? max(0, min(TC_M, p.M - m0)) : min(TC_M, p.M);
  const int nchunks = (kend - kbeg + TC_KC - 1) / TC_KC;
  const int kend_b = min(kend, Kb);

  auto issue_chunk = [&](int c) {
    if (c < nchunks) {
      const int slot = c % RS, k0 = kbeg + c * TC_KC;
      raw_issue<AMAJ, TC_M>(rawA(slot), Abase, p.lda, TC_M, a_rows_valid, k0, kend);
      if (BPREP) {
        const float4* img = bt.Bprep + (size_t)(k0 / TC_KC) * (2 * b_f4);
        float4* dst = sBhi(0, c);
        const int total = 2 * (int)b_f4;
#pragma unroll
        for (int q = tid; q < total; q += TC_THREADS) cp_async16_zfill(dst + q, img + q, 16);
      } else {
        raw_issue<BMAJ, NT_T>(rawB(slot), bt.B, p.ldb, NT, bt.n_valid, k0, kend_b);
      }
    }
    cp_async_commit_group();  // always commit so the group count stays uniform
  };
  for (int c = 0; c < RS; ++c) issue_chunk(c);  // global loads start before the TMEM allocation settles

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  const uint32_t idesc = umma_idesc_tf32(TC_M, NT);
  const uint32_t a_lbo = TC_M * 16u, b_lbo = (uint32_t)NT * 16u, sbo = 128u;
  const uint32_t a_kstep = 2u * TC_M * 16u, b_kstep = 2u * (uint32_t)NT * 16u;  // bytes per UMMA k-step (8 tf32)

  for (int c = 0; c < nchunks; ++c) {
    if (RS == 3) cp_async_wait_group<2>(); else cp_async_wait_group<1>();   // this thread's pieces of chunk c landed
    __syncthreads();                                                         // ... and everybody else's
    const int s = c & 1, slot = c % RS, k0 = kbeg + c * TC_KC;
    if (c >= 2) mbar_wait(&s_bar[s], (uint32_t)(((c >> 1) - 1) & 1));        // MMAs of chunk c-2 released stage s
    raw_convert<AMAJ, TC_M>(rawA(slot), TC_M, k0, kend, -1, sAhi(s), sAlo(s));
    if (!BPREP)
      raw_convert<BMAJ, NT_T>(rawB(slot), NT, k0, kend_b, BMAJ == TC_MN_MAJOR ? p.ones_col : -1, sBhi(s, c), sBlo(s, c));
    fence_async_smem();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const uint32_t ah = smem_u32(sAhi(s)), al = smem_u32(sAlo(s));
      const uint32_t bh = smem_u32(sBhi(s, c)), bl = smem_u32(sBlo(s, c));
#pragma unroll
      for (int ks = 0; ks < TC_KC / 8; ++ks) {
        const uint64_t dah = umma_desc(ah + ks * a_kstep, a_lbo, sbo);
        const uint64_t dal = umma_desc(al + ks * a_kstep, a_lbo, sbo);
        const uint64_t dbh = umma_desc(bh + ks * b_kstep, b_lbo, sbo);
        const uint64_t dbl = umma_desc(bl + ks * b_kstep, b_lbo, sbo);
        const uint32_t first = (c == 0 && ks == 0) ? 0u : 1u;
        if (p.dbg & 2) {
          umma_tf32(tmem, dah, dbh, idesc, first);
        } else {
          umma_tf32(tmem, dal, dbh, idesc, first);  // small terms first
          umma_tf32(tmem, dah, dbl, idesc, 1u);
          umma_tf32(tmem, dah, dbh, idesc, 1u);
        }
      }
      umma_commit(&s_bar[s]);
    }
    // Refill: raw slot `slot` was fully consumed before the barrier above. With BPREP the B image
    // of chunk c+2 lands in the ring slot of chunk c-1, whose MMAs (issued one iteration ago) must be done.
    if (BPREP && c >= 1 && c + RS < nchunks) mbar_wait(&s_bar[(c - 1) & 1], (uint32_t)(((c - 1) >> 1) & 1));
    issue_chunk(c + RS);
  }
  cp_async_wait_group<0>();
  if (nchunks > 0) {
    const int last = nchunks - 1;
    mbar_wait(&s_bar[last & 1], (uint32_t)((last >> 1) & 1));
  }
  tc_fence_after();

  // ---- epilogue: warp w reads TMEM lanes 32*(w%4).., column groups of 16 alternate between warp halves ----
  const int row = (warp & 3) * 32 + lane;
  const int half = warp >> 2;
  const uint32_t lane_addr = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  const long grow = (long)m0 + row;
  const bool row_ok = row < a_rows_valid;
#pragma unroll 1
  for (int n0 = half * 16; n0 < NT; n0 += 32) {
    float v[16];
    if (nchunks > 0) {
      tmem_ld16(lane_addr + (uint32_t)n0, v);
    } else {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = 0.f;
    }
    if (row_ok) tc_epilogue_row<AMAJ>(p, bt, v, n0, row, grow);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, ncols);
}

}  // namespace ed3dgs
