// The lattice-sweep engine: one persistent CTA walks a tile of configurations through every lattice
// site.  Per site it (1) gathers the neighbours' hidden states from an L2-resident history ring,
// (2) runs the gate GEMM with FP32 FFMA against weights streamed from L2 through a double-buffered
// shared-memory stage (class-sorted rows for the tensorized cells, whose weight slice depends on the
// neighbouring spins), (3) applies the cell non-linearity in registers, and (4) evaluates the softmax /
// phase heads with warp shuffles, the U(1) mask, the inverse-CDF draw (sampling) or the log-probability
// gather (evaluation), accumulating log psi per row in float64.
//
// Reference behaviour restated here (file:line under the reference root):
//   1DTFIM/RNNwavefunction.py:35-118, J1J2/ComplexRNNwavefunction.py:45-169,
//   2DTFIM_2DRNN/RNNwavefunction.py:35-200 + MDRNNcell.py:51-66,
//   2DHeisenberg/MDComplexRNN.py:69-286 + MDTensorizedcell.py:40-49 + MDTensorizedGRUcell.py:49-61,
//   RydbergKagome_2DGRU/MDRNNWavefunction.py:73-238 + MDGRUcell.py:48-59.
#pragma once
#include "common.cuh"

namespace rnnwf {

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src));
}
// 4-byte variant with zero-fill: src_bytes == 0 writes zeros without touching global memory
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gmem_src, int src_bytes) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(s), "l"(gmem_src), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }

__device__ __forceinline__ float sigmoid_f(float x) { return 1.0f / (1.0f + expf(-x)); }

template <int TJ, int NTU>
__global__ void __launch_bounds__(MAX_THREADS, 1) sweep_kernel(const SweepParams P) {
  extern __shared__ __align__(16) unsigned char smem[];
  const SmemLayout L(P.ncls_max, P.Gblk, P.Mp, P.ldH, P.nhp, P.ND, P.Tsym);
  float4* const Ws = reinterpret_cast<float4*>(smem + L.off_W);
  float* const As = reinterpret_cast<float*>(smem + L.off_A);
  float* const Hs = reinterpret_cast<float*>(smem + L.off_H);
  float* const headW = reinterpret_cast<float*>(smem + L.off_headw);
  float* const headB = reinterpret_cast<float*>(smem + L.off_headb);
  uint8_t* const symS = reinterpret_cast<uint8_t*>(smem + L.off_sym);
  double* const rlogp = reinterpret_cast<double*>(smem + L.off_logp);
  double* const rphase = reinterpret_cast<double*>(smem + L.off_phase);
  int* const rnup = reinterpret_cast<int*>(smem + L.off_nup);
  int* const rparent = reinterpret_cast<int*>(smem + L.off_parent);
  float* const rcoef = reinterpret_cast<float*>(smem + L.off_coef);
  int* const rpos = reinterpret_cast<int*>(smem + L.off_rpos);
  int* const cnt = reinterpret_cast<int*>(smem + L.off_cnt);  // [0..7] counts, [8..15] bases
  int16_t* const perm_s = reinterpret_cast<int16_t*>(smem + L.off_perm);
  int16_t* const perm_id = reinterpret_cast<int16_t*>(smem + L.off_permid);
  uint8_t* const qcls_s = reinterpret_cast<uint8_t*>(smem + L.off_qcls);
  uint8_t* const qzero = reinterpret_cast<uint8_t*>(smem + L.off_qzero);
  uint8_t* const rcls = reinterpret_cast<uint8_t*>(smem + L.off_rcls);

  const int tid = threadIdx.x, nthreads = blockDim.x;
  const int warp = tid >> 5, lane = tid & 31, nwarps = nthreads >> 5;
  const int wr = warp / P.WC, wc = warp - wr * P.WC;
  const int rl = lane >> 2, cl = lane & 3;
  const int p0 = (wr * RL + rl) * TM;     // first of this thread's 4 consecutive tile positions
  const int gcol = wc * CL + cl;          // this thread's column groups: g_i = i*gstride + gcol
  const int gstride = CL * P.WC;
  const int Mp = P.Mp, ldA = Mp + 4, ldH = P.ldH, nhp = P.nhp;
  const int wstage = P.ncls_max * KC * P.Gblk;
  float* const ring = P.ring + (size_t)blockIdx.x * P.ring_len * Mp * nhp;

  for (int i = tid; i < nhp * P.ND; i += nthreads) headW[i] = P.head_w[i];
  if (tid < P.ND) headB[tid] = P.head_b[tid];
  for (int i = tid; i < Mp / 4; i += nthreads) qzero[i] = 0;

  const long long total_rows = (P.rows_dev != nullptr) ? (long long)(*P.rows_dev) : P.rows;
  const long long ntiles = (total_rows + P.Mr - 1) / P.Mr;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long row0 = tile * P.Mr;
    const int nrows = (int)((total_rows - row0 < (long long)P.Mr) ? (total_rows - row0) : (long long)P.Mr);
    // prefix reuse: children are ordered by their first changed symbol, so the tile's first child has
    // the smallest one; every row of the tile restarts there from its parent's saved history.
    int ts0 = 0;
    if (P.mode == MODE_CHILD && P.par_hist != nullptr) {
      const Child c0 = P.children[row0];
      ts0 = (c0.tb >= 0 && c0.tb < c0.ta) ? c0.tb : c0.ta;
    }
    const int q0 = ts0 * P.symL;
    __syncthreads();  // the previous tile is completely done with shared memory

    for (int i = tid; i < Mp; i += nthreads) {
      perm_id[i] = (int16_t)(i < nrows ? i : -1);
      rlogp[i] = 0.0;
      rphase[i] = 0.0;
      rnup[i] = 0;
      rparent[i] = 0;
      rcoef[i] = 0.f;
    }
    for (int i = tid; i < Mp * ldH; i += nthreads) Hs[i] = 0.f;
    if (P.mode != MODE_SAMPLE) {
      for (int idx = tid; idx < nrows * P.Tsym; idx += nthreads) {
        const int r = idx / P.Tsym, t = idx - r * P.Tsym;
        const long long src = (P.mode == MODE_CHILD) ? (long long)P.children[row0 + r].parent : row0 + r;
        const int site = P.steps[t * P.symL + P.symL - 1].site;
        const int8_t* c = P.cfg_in + src * P.nsites;
        int sym;
        if (P.bits == 1) sym = c[site] & 1;
        else sym = (c[site * 3] & 1) | ((c[site * 3 + 1] & 1) << 1) | ((c[site * 3 + 2] & 1) << 2);
        symS[t * Mp + r] = (uint8_t)sym;
      }
    }
    __syncthreads();
    if (P.mode == MODE_CHILD && tid < nrows) {
      const Child ch = P.children[row0 + tid];
      symS[ch.ta * Mp + tid] ^= (uint8_t)(1u << ch.ka);
      if (ch.tb >= 0) symS[ch.tb * Mp + tid] ^= (uint8_t)(1u << ch.kb);
      rparent[tid] = ch.parent;
      rcoef[tid] = ch.coef;
      int nup = 0;
      for (int t = 0; t < ts0; ++t) nup += symS[t * Mp + tid];
      rnup[tid] = nup;
    }
    __syncthreads();

    auto hist_ptr = [&](int qn, int r) -> const float* {
      if (P.par_hist != nullptr && qn < q0)
        return P.par_hist + ((size_t)qn * (size_t)P.par_rows + (size_t)rparent[r]) * nhp;
      return ring + ((size_t)(qn % P.ring_len) * Mp + r) * nhp;
    };

    for (int q = q0; q < P.Q; ++q) {
      const StepMeta* sm = P.steps + q;
      const VariantMeta* vm = P.variants + sm->variant;
      const int ncls = vm->ncls;
      // active GEMM slots of this step, packed 4 bits each
      int apack = 0, nact = 0;
#pragma unroll
      for (int s = 0; s < MAX_SLOTS; ++s)
        if (((vm->slot_mask >> s) & 1) && sm->nbr[s] >= 0) { apack |= s << (4 * nact); ++nact; }
      const int nchunks = nact * P.cps;

      // ---- class sort (tensorized cells): rows with the same (left, up) inputs share a weight slice
      const int16_t* perm = perm_id;
      const uint8_t* qcls = qzero;
      if (ncls > 1) {
        if (tid < 16) cnt[tid] = 0;
        for (int i = tid; i < Mp; i += nthreads) perm_s[i] = -1;
        __syncthreads();
        if (tid < nrows) {
          int c = 0;
#pragma unroll
          for (int s = 0; s < MAX_SLOTS; ++s)
            if (vm->cls_mult[s] != 0 && sm->nsym[s] >= 0) c += vm->cls_mult[s] * symS[sm->nsym[s] * Mp + tid];
          rcls[tid] = (uint8_t)c;
          rpos[tid] = atomicAdd(&cnt[c], 1);
        }
        __syncthreads();
        if (tid == 0) {
          int off = 0;
          for (int c = 0; c < ncls; ++c) { cnt[8 + c] = off; off += (cnt[c] + 3) & ~3; }
        }
        __syncthreads();
        if (tid < nrows) perm_s[cnt[8 + rcls[tid]] + rpos[tid]] = (int16_t)tid;
        for (int qd = tid; qd < Mp / 4; qd += nthreads) {
          int c = 0;
          for (int k = 1; k < ncls; ++k) if (cnt[8 + k] <= qd * 4) c = k;
          qcls_s[qd] = (uint8_t)c;
        }
        __syncthreads();
        perm = perm_s;
        qcls = qcls_s;
      }
      const int mycls = qcls[p0 >> 2];

      for (int cb = 0; cb < P.ncb; ++cb) {
        if (cb > 0) __syncthreads();  // the previous column block's last stage is still being read
        float acc[TM][TJ][NTU];
#pragma unroll
        for (int a = 0; a < TM; ++a)
#pragma unroll
          for (int b = 0; b < TJ; ++b)
#pragma unroll
            for (int c = 0; c < NTU; ++c) acc[a][b][c] = 0.f;

        if (nchunks > 0) {
          // One pipeline stage = the class-sliced weights of one k-chunk (bulk 16 B cp.async) plus the matching
          // 8 columns of the neighbours' hidden states, transposed on the fly into As[kk][position] by 4 B cp.async
          // (zero-filled for padding positions) -- nothing is staged through registers.
          auto issue = [&](int ci, int buf) {
            const int sidx = ci / P.cps, c = ci - sidx * P.cps;
            const int s = (apack >> (4 * sidx)) & 15;
            const int n4 = ncls * KC * P.Gblk;
            const float4* src = P.Wpack + vm->w_off + (size_t)cb * vm->w_cb_stride +
                                (size_t)(vm->slot_rank[s] * P.cps + c) * n4;
            float4* dst = Ws + buf * wstage;
            for (int i = tid; i < n4; i += nthreads) cp_async16(dst + i, src + i);
            const int qn = sm->nbr[s];
            float* abase = As + buf * KC * ldA;
            for (int idx = tid; idx < Mp * KC; idx += nthreads) {
              const int pos = idx >> 3, kk = idx & 7;
              const int r = perm[pos];
              const float* g = (r >= 0) ? hist_ptr(qn, r) + c * KC + kk : P.head_b;
              cp_async4(abase + kk * ldA + pos, g, r >= 0 ? 4 : 0);
            }
            cp_async_commit();
          };
          issue(0, 0);
          for (int ci = 0; ci < nchunks; ++ci) {
            const int buf = ci & 1;
            cp_async_wait_all();
            __syncthreads();
            if (ci + 1 < nchunks) issue(ci + 1, buf ^ 1);
            const float* Ab = As + buf * KC * ldA + p0;
            const float4* Wb = Ws + buf * wstage + mycls * KC * P.Gblk + gcol;
#pragma unroll
            for (int kk = 0; kk < KC; ++kk) {
              const float4 a = *reinterpret_cast<const float4*>(Ab + kk * ldA);
#pragma unroll
              for (int i = 0; i < TJ; ++i) {
                const float4 w = Wb[kk * P.Gblk + i * gstride];
                acc[0][i][0] = fmaf(a.x, w.x, acc[0][i][0]); acc[1][i][0] = fmaf(a.y, w.x, acc[1][i][0]);
                acc[2][i][0] = fmaf(a.z, w.x, acc[2][i][0]); acc[3][i][0] = fmaf(a.w, w.x, acc[3][i][0]);
                acc[0][i][1] = fmaf(a.x, w.y, acc[0][i][1]); acc[1][i][1] = fmaf(a.y, w.y, acc[1][i][1]);
                acc[2][i][1] = fmaf(a.z, w.y, acc[2][i][1]); acc[3][i][1] = fmaf(a.w, w.y, acc[3][i][1]);
                acc[0][i][2] = fmaf(a.x, w.z, acc[0][i][2]); acc[1][i][2] = fmaf(a.y, w.z, acc[1][i][2]);
                acc[2][i][2] = fmaf(a.z, w.z, acc[2][i][2]); acc[3][i][2] = fmaf(a.w, w.z, acc[3][i][2]);
                if (NTU == 4) {
                  acc[0][i][NTU - 1] = fmaf(a.x, w.w, acc[0][i][NTU - 1]); acc[1][i][NTU - 1] = fmaf(a.y, w.w, acc[1][i][NTU - 1]);
                  acc[2][i][NTU - 1] = fmaf(a.z, w.w, acc[2][i][NTU - 1]); acc[3][i][NTU - 1] = fmaf(a.w, w.w, acc[3][i][NTU - 1]);
                }
              }
            }
          }
        }

        // ---- epilogue: cell non-linearity, thread-local
#pragma unroll
        for (int i = 0; i < TJ; ++i) {
          const int g = cb * P.Gblk + i * gstride + gcol;
          if (g >= P.Greal) continue;
          const float4 b4 = __ldg(P.bias + vm->b_off + g);
#pragma unroll
          for (int rr = 0; rr < TM; ++rr) {
            const int r = perm[p0 + rr];
            if (r < 0) continue;
            float4 x4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (vm->x_mask) {
#pragma unroll
              for (int s = 0; s < MAX_SLOTS; ++s) {
                if (((vm->x_mask >> s) & 1) && sm->nsym[s] >= 0) {
                  const int sy = symS[sm->nsym[s] * Mp + r];
                  const float4 t4 = __ldg(P.Xtab + vm->x_off + (size_t)(s * P.D + sy) * P.Gp + g);
                  x4.x += t4.x; x4.y += t4.y; x4.z += t4.z; x4.w += t4.w;
                }
              }
            }
            float4 gate = make_float4(0.f, 0.f, 0.f, 0.f);
            if (P.cell == CELL_TENSGRU2D || P.cell == CELL_GRU4) {
              const float u = sigmoid_f(acc[rr][i][1] + b4.y + x4.y);
              const float ht = tanhf(acc[rr][i][0] + b4.x + x4.x);
              const float m = acc[rr][i][2];
              Hs[r * ldH + g] = u * ht + (1.f - u) * m;
              gate = make_float4(u, ht, m, 0.f);
            } else if (P.cell == CELL_GRU1D) {
              const float rg = sigmoid_f(acc[rr][i][0] + b4.x + x4.x);
              const float u = sigmoid_f(acc[rr][i][1] + b4.y + x4.y);
              const float hh = acc[rr][i][NTU - 1] + b4.w;
              const float c = tanhf(acc[rr][i][2] + b4.z + x4.z + rg * hh);
              float hown = 0.f;
              if (sm->own != 0xff && sm->nbr[sm->own] >= 0) hown = __ldcg(hist_ptr(sm->nbr[sm->own], r) + g);
              Hs[r * ldH + g] = (1.f - u) * c + u * hown;
              gate = make_float4(rg, u, c, hh);
            } else {
              float v[4] = {acc[rr][i][0] + b4.x + x4.x, acc[rr][i][1] + b4.y + x4.y, acc[rr][i][2] + b4.z + x4.z,
                            acc[rr][i][NTU - 1] + b4.w + x4.w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float hv = (P.cell == CELL_VANILLA2D) ? (v[e] > 0.f ? v[e] : expm1f(v[e])) : tanhf(v[e]);
                if (4 * g + e < nhp) Hs[r * ldH + 4 * g + e] = hv;
              }
            }
            if (P.sv_gates != nullptr && P.mode != MODE_CHILD)
              P.sv_gates[((size_t)q * (size_t)P.sv_rows + (size_t)(row0 + r)) * P.Gp + g] = gate;
          }
        }
      }  // column blocks
      __syncthreads();

      // ---- publish the new hidden states: history ring (+ saved history for backward / prefix reuse)
      {
        float* dst = ring + (size_t)(q % P.ring_len) * Mp * nhp;
        const int n4 = nhp >> 2;
        for (int idx = tid; idx < nrows * n4; idx += nthreads) {
          const int r = idx / n4, c4 = idx - r * n4;
          const float4 v = *reinterpret_cast<const float4*>(Hs + r * ldH + c4 * 4);
          *reinterpret_cast<float4*>(dst + (size_t)r * nhp + c4 * 4) = v;
          if (P.sv_hist != nullptr && P.mode != MODE_CHILD)
            *reinterpret_cast<float4*>(P.sv_hist + ((size_t)q * (size_t)P.sv_rows + (size_t)(row0 + r)) * nhp + c4 * 4) = v;
        }
      }

      // ---- output heads
      if (sm->head) {
        const int t = sm->sym;
        const int grp = lane >> 3, l8 = lane & 7;
        for (int rb = warp * 4; rb < Mp; rb += nwarps * 4) {
          const int r = rb + grp;
          const bool valid = r < nrows;
          float part[8];
#pragma unroll
          for (int d = 0; d < 8; ++d) part[d] = 0.f;
          if (valid) {
            const float* hrow = Hs + r * ldH;
            for (int j = l8; j < P.nh; j += HL) {
              const float hv = hrow[j];
              const float* w = headW + j * P.ND;
#pragma unroll
              for (int d = 0; d < 8; ++d)
                if (d < P.ND) part[d] = fmaf(hv, w[d], part[d]);
            }
          }
#pragma unroll
          for (int d = 0; d < 8; ++d) {
            if (d < P.ND) {
              part[d] += __shfl_xor_sync(0xffffffffu, part[d], 1);
              part[d] += __shfl_xor_sync(0xffffffffu, part[d], 2);
              part[d] += __shfl_xor_sync(0xffffffffu, part[d], 4);
            }
          }
          if (valid && l8 == 0) {
            float p[8];
            float mx = -INFINITY;
#pragma unroll
            for (int d = 0; d < 8; ++d)
              if (d < P.D) { p[d] = part[d] + headB[d]; mx = fmaxf(mx, p[d]); } else p[d] = 0.f;
            float ssum = 0.f;
#pragma unroll
            for (int d = 0; d < 8; ++d)
              if (d < P.D) { p[d] = expf(p[d] - mx); ssum += p[d]; }
#pragma unroll
            for (int d = 0; d < 8; ++d) p[d] = p[d] / ssum;
            int nup = 0;
            if (P.u1) {  // U(1) projection: forbid the species that is already exhausted, renormalise (L1)
              nup = rnup[r];
              const int ndn = t - nup;
              p[0] *= (P.cap_dn - ndn >= 0) ? 1.f : 0.f;
              p[1] *= (P.cap_up - nup >= 0) ? 1.f : 0.f;
              const float nrm = p[0] + p[1];
              p[0] /= nrm;
              p[1] /= nrm;
            }
            int s;
            if (P.mode == MODE_SAMPLE) {
              const float u = (P.uniforms != nullptr)
                                  ? P.uniforms[(size_t)(row0 + r) * P.Tsym + t]
                                  : philox_uniform(P.seed, P.row_offset + (unsigned long long)(row0 + r), (uint32_t)t);
              double tot = 0.0;
#pragma unroll
              for (int d = 0; d < 8; ++d) if (d < P.D) tot += (double)p[d];
              const double target = (double)u * tot;
              double c = 0.0;
              s = 0;
#pragma unroll
              for (int d = 0; d < 8; ++d)
                if (d < P.D) { c += (double)p[d]; if (c <= target) ++s; }
              if (s > P.D - 1) s = P.D - 1;
              symS[t * Mp + r] = (uint8_t)s;
            } else {
              s = symS[t * Mp + r];
            }
            float ps = 0.f, xs = 0.f;
#pragma unroll
            for (int d = 0; d < 8; ++d) if (d == s) ps = p[d];
            const double lp = rlogp[r] + log((double)ps);
            rlogp[r] = lp;
            double ph = rphase[r];
            if (P.complex_head) {
              xs = (s == 0) ? (part[2] + headB[2]) : (part[3] + headB[3]);  // complex models are two-state (D == 2)
              ph += 3.14159265358979323846 * ((double)xs / (1.0 + fabs((double)xs)));
              rphase[r] = ph;
            }
            if (P.u1) rnup[r] = nup + s;
            if (P.sv_cum != nullptr && P.mode != MODE_CHILD)
              P.sv_cum[(size_t)t * (size_t)P.sv_rows + (size_t)(row0 + r)] = make_double2(lp, ph);
          }
        }
      }
      __syncthreads();
    }  // steps

    // ---- tile results
    if (P.mode == MODE_CHILD) {
      if (tid < nrows) {
        const int par = rparent[tid];
        double base_lp = P.par_logp[par], base_ph = (P.par_phase != nullptr) ? P.par_phase[par] : 0.0;
        if (ts0 > 0 && P.par_cum != nullptr) {
          const double2 c = P.par_cum[(size_t)(ts0 - 1) * (size_t)P.par_rows + par];
          base_lp -= c.x;
          base_ph -= c.y;
        }
        const double amp = exp(0.5 * (rlogp[tid] - base_lp)) * (double)rcoef[tid];
        if (P.complex_head) {
          double sn, cs;
          sincos(rphase[tid] - base_ph, &sn, &cs);
          atomicAdd(P.eloc + 2 * (size_t)par, amp * cs);
          atomicAdd(P.eloc + 2 * (size_t)par + 1, amp * sn);
        } else {
          atomicAdd(P.eloc + 2 * (size_t)par, amp);
        }
      }
    } else {
      if (tid < nrows) {
        if (P.out_logp != nullptr) P.out_logp[row0 + tid] = rlogp[tid];
        if (P.out_phase != nullptr) P.out_phase[row0 + tid] = rphase[tid];
      }
      if (P.mode == MODE_SAMPLE && P.cfg_out != nullptr) {
        for (int idx = tid; idx < nrows * P.Tsym; idx += nthreads) {
          const int r = idx / P.Tsym, t = idx - r * P.Tsym;
          const int site = P.steps[t * P.symL + P.symL - 1].site;
          const int sy = symS[t * Mp + r];
          int8_t* c = P.cfg_out + (size_t)(row0 + r) * P.nsites;
          if (P.bits == 1) c[site] = (int8_t)sy;
          else { c[site * 3] = (int8_t)(sy & 1); c[site * 3 + 1] = (int8_t)((sy >> 1) & 1); c[site * 3 + 2] = (int8_t)((sy >> 2) & 1); }
        }
      }
    }
  }  // tiles
}

}  // namespace rnnwf
