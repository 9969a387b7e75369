// hf2d_stage1.cuh -- the ordered part of Stage 1 of the DEEPS2D step (deeps2d_core.cpp:853-1026): the Neumann
// in-place overwrite as a parallel pre-pass, and the per-step scalar kernels.  Stage 1 proper lives in k_point
// (hf2d_point.cuh).
//
// The reference sweeps i-outer/j-inner IN PLACE: a node with a d/dx=0 or d/dy=0 flag overwrites its own S[k]
// from its neighbours before the update, and nodes visited later see the overwritten value while nodes visited
// earlier saw the old one.  Parallel-exact restatement:
//   k_neumann_save   : copy S of every Neumann-list node into its SpecNode::S0            (old values)
//   k_neumann_level  : level by level (level = longest chain of Neumann nodes through left/down neighbours)
//                      apply the overwrite in place, reading left/down from the array (already new) and
//                      right/up through S0 when that neighbour is itself a list node (still old)
//   k_point (Stage 1): every node: own/left/down S from the array (new), right/up S through S0 if listed (old)
#pragma once
#include "hf2d_dev.cuh"

__global__ void k_begin_step(Params P) {
    // one thread: dt for this step, scenario tables, FillNode2D mode flags; clears the residual accumulators
    StepState* s = P.st;
    const double dtm = bits2d(s->dt_min_bits);
    s->dt = dtm;                                                     // deeps2d_core.cpp:803 (serial rule)
    if (P.dt_rule == 1) s->dt_min_bits = d2bits(1.0);               // MPI rule: per-step minimum (:850)
    const double it = (double)s->iter;
    DivExact dv;
    s->beta_scn = tab_get(dv, &P.tab[13], it);
    s->CFL_scn = tab_get(dv, &P.tab[12], it);
    if ((int)s->iter < P.TurbStartIter) { s->is_mu_t = 0; s->is_init = P.isTurbulenceReset; }
    else { s->is_mu_t = 1; s->is_init = 0; }
    for (int k = 0; k < NEQ; k++) { s->RMS[k] = 0.; s->sumDiv[k] = 0.; s->iRMS[k] = 0ull; s->DDmax_bits[k] = 0ull; s->nan_last[k] = 0ull; s->DDfix_bits[k] = 0ull; }
    s->block_counter = 0;
}

__global__ void k_end_step(Params P) {
    StepState* s = P.st;
    s->time_part += s->dt;                                           // CurrentTimePart += dt, :1602
    s->iter += 1;
}

__global__ void k_neumann_save(Params P, const int* __restrict__ list, int n_list) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_list) return;
    SpecNode& sp = P.spec[list[t]];
    const long long n = sp.node;
#pragma unroll
    for (int k = 0; k < NEQ; k++) sp.S0[k] = P.S_in[(long long)k * P.N + n];
}

// S of neighbour m as a node AHEAD of it in the sweep sees it: the pre-overwrite value
__device__ __forceinline__ double s_old(const Params& P, long long m, uint32_t tt_m, int k) {
    if (tt_m & TT_NLIST) return P.spec[P.spec_idx[m]].S0[k];
    return P.S_in[(long long)k * P.N + m];
}

__global__ void k_neumann_level(Params P, const int* __restrict__ list, int first, int count) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const SpecNode& sp = P.spec[list[first + t]];
    const long long n = sp.node, ny = P.ny, N = P.N;
    const uint32_t ct = P.ct[n], tt = P.tt[n];
    const int n1 = (tt & TT_IDXL) ? 1 : 0, n2 = (tt & TT_IDXR) ? 1 : 0, n3 = (tt & TT_IDYU) ? 1 : 0, n4 = (tt & TT_IDYD) ? 1 : 0;
    const long long nL = n - n1 * ny, nR = n + n2 * ny, nU = n + n3, nD = n - n4;
    const double n_n = OMAX(n1 + n2, 1), m_m = OMAX(n3 + n4, 1);
    const double n_n_1 = 1. / n_n, m_m_1 = 1. / m_m;
    const uint32_t ttR = P.tt[nR], ttU = P.tt[nU];
    const int ne = num_eq(tt);
    for (int k = 0; k < ne; k++) {
        const unsigned f = eq_flags(ct, tt, k, P.ns);
        if (!(f & 1)) continue;
        double* Sk = P.S_in + (long long)k * N;
        double sc = sp.S0[k];      // the node's own pre-overwrite value (k_neumann_save): re-evaluation after a hand-off is idempotent
        bool wr = false;
        if (!(f & 2)) {   // :998  (weights are crossed in x: left*n2 + right*n1)
            const double sl = n1 ? Sk[nL] : sc;
            const double sr = n2 ? s_old(P, nR, ttR, k) : sc;
            sc = (sl * n2 + sr * n1) * n_n_1;
            wr = true;
        }
        if (!(f & 4)) {   // :1005
            const double su = n3 ? s_old(P, nU, ttU, k) : sc;
            const double sd = n4 ? Sk[nD] : sc;
            sc = (su * n3 + sd * n4) * m_m_1;
            wr = true;
        }
        if (wr) Sk[n] = sc;
    }
}

// Chains: when an overwritten S[k] of a node depends on the overwritten S[k] of its left/down neighbour (e.g. the
// eps equation along a whole d/dy=0 row, a side effect of the flag quirk at deeps2d_core.cpp:918-930) the values
// form a first-order recurrence that must be evaluated in sweep order to stay bit-exact.  k_neumann_level has
// already produced every link-free value; the chain kernel then walks the links, carrying the running value in a
// register.  (A first version fetched 32 links per warp iteration: two dependent global-load latencies per 32 links,
// 117 us for a 2048-link row; staging whole chunks first removes them from the walk.)
#define CHAIN_THREADS 256
#define CHAIN_CHUNK   1024      // links staged per pass: 4 per thread
__global__ void __launch_bounds__(CHAIN_THREADS) k_neumann_chains(Params P, const int* __restrict__ chain_start, const int* __restrict__ chain_src,
                                                                  const ChainElem* __restrict__ el, int n_chains) {
    // One BLOCK per chain.  All threads stage CHAIN_CHUNK links at once (link record, then its `other` operand: the two
    // dependent global loads are in flight for the whole chunk together), warp 0 replays the recurrence from shared
    // memory 32 links at a time (broadcasts issue ahead of the carry: critical path mul -> add -> mul per link), then all
    // threads write the results.  None of the operands is produced by a chain (hf2d_io.inc), so staging ahead is safe.
    __shared__ double s_t[CHAIN_CHUNK];
    __shared__ float2 s_w[CHAIN_CHUNK];
    const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    if (c >= n_chains) return;
    const int e0 = chain_start[c], e1 = chain_start[c + 1];
    const long long N = P.N;
    double carry = 0.;
    if (tid < 32) carry = P.S_in[(long long)el[e0].k * N + chain_src[c]];
    for (int eb = e0; eb < e1; eb += CHAIN_CHUNK) {
        const int cnt = min(CHAIN_CHUNK, e1 - eb);
        int node[CHAIN_CHUNK / CHAIN_THREADS]; short kk[CHAIN_CHUNK / CHAIN_THREADS];
#pragma unroll
        for (int m = 0; m < CHAIN_CHUNK / CHAIN_THREADS; m++) {
            const int i = tid + m * CHAIN_THREADS;
            node[m] = -1; kk[m] = 0;
            if (i < cnt) {
                const ChainElem q = el[eb + i];
                const double o = q.other_spec ? P.spec[q.other].S0[q.k] : P.S_in[(long long)q.k * N + q.other];
                s_t[i] = o * (double)q.wo;
                s_w[i] = make_float2(q.wc, q.sc);
                node[m] = q.node; kk[m] = q.k;
            }
        }
        __syncthreads();
        if (tid < 32) {
            for (int b0 = 0; b0 < cnt; b0 += 32) {
                const int n32 = min(32, cnt - b0);
                double t = 0., wcd = 0., scd = 0.;
                if (lane < n32) { t = s_t[b0 + lane]; const float2 w = s_w[b0 + lane]; wcd = (double)w.x; scd = (double)w.y; }
                double mine = 0.;
                // weight 1 on the carried value (the usual case: both x- or y-neighbours present): x * 1.0 == x exactly,
                // so that multiplication leaves the dependent chain
                const bool w1 = __all_sync(0xffffffffu, wcd == 1.0);
                if (n32 == 32 && w1) {
#pragma unroll
                    for (int q = 0; q < 32; q++) {
                        const double ts = __shfl_sync(0xffffffffu, t, q);
                        const double sc = __shfl_sync(0xffffffffu, scd, q);
                        carry = (carry + ts) * sc;
                        if (q == lane) mine = carry;
                    }
                } else if (n32 == 32) {
#pragma unroll
                    for (int q = 0; q < 32; q++) {
                        const double ts = __shfl_sync(0xffffffffu, t, q);
                        const double wc = __shfl_sync(0xffffffffu, wcd, q), sc = __shfl_sync(0xffffffffu, scd, q);
                        carry = (carry * wc + ts) * sc;
                        if (q == lane) mine = carry;
                    }
                } else {
                    for (int q = 0; q < n32; q++) {
                        const double ts = __shfl_sync(0xffffffffu, t, q);
                        const double wc = __shfl_sync(0xffffffffu, wcd, q), sc = __shfl_sync(0xffffffffu, scd, q);
                        carry = (carry * wc + ts) * sc;
                        if (q == lane) mine = carry;
                    }
                }
                if (lane < n32) s_t[b0 + lane] = mine;
            }
        }
        __syncthreads();
#pragma unroll
        for (int m = 0; m < CHAIN_CHUNK / CHAIN_THREADS; m++)
            if (node[m] >= 0) P.S_in[(long long)kk[m] * N + node[m]] = s_t[tid + m * CHAIN_THREADS];
        __syncthreads();
    }
}

// this-sweep value of dSdx[k] (dir 0) / dSdy[k] (dir 1) of node m, for soft-BC readers behind it in the sweep
template <class KC>
__device__ __forceinline__ double ds_new(const Params& P, long long m, int k, int dir) {
    const uint32_t ct = P.ct[m], tt = P.tt[m];
    const double* oldp = dir ? P.dSy_in[k] : P.dSx_in[k];
    if (!node_active(ct) || k >= num_eq(tt)) return oldp[m];
    const unsigned f = eq_flags(ct, tt, k, KC::ns);
    if (!(f & 1)) return oldp[m];
    if (dir == 0) {
        if (!(f & 2)) return 0.;
        const int n1 = (tt & TT_IDXL) ? 1 : 0, n2 = (tt & TT_IDXR) ? 1 : 0;
        const double n_n_1 = 1. / (double)OMAX(n1 + n2, 1);
        const double* Ak = P.A + (long long)k * P.N;
        return (Ak[m + n2 * (long long)P.ny] - Ak[m - n1 * (long long)P.ny]) * n_n_1;
    } else {
        if (!(f & 4)) return 0.;
        const int n3 = (tt & TT_IDYU) ? 1 : 0, n4 = (tt & TT_IDYD) ? 1 : 0;
        const double m_m_1 = 1. / (double)OMAX(n3 + n4, 1);
        const double* Bk = P.B + (long long)k * P.N;
        return (Bk[m + n3] - Bk[m - n4]) * m_m_1;
    }
}
