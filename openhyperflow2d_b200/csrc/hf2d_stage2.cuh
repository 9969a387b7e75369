// hf2d_stage2.cuh -- Stage 2 of the DEEPS2D step (deeps2d_core.cpp:1028-1334) as ONE kernel family:
// residual + blending factor, commit, mixed old/new gradients, FillNode2D, TurbModRANS2D, CFL dt, chemistry.
//
// Work decomposition ("strip sweep").  A warp owns a strip of 31 consecutive rows (y is the contiguous index, so a
// warp's loads are one 256-byte segment per plane) and walks TI columns in +x, which is the reference's outer sweep
// direction.  In the reference's in-place sweep node C reads its LEFT and DOWN neighbours after they went through
// commit + FillNode2D + chemistry in this step and its RIGHT and UP neighbours before.  What a
// neighbour exposes (S[0], S[4..8], U, V, Tg) depends only on its own shadow S, flags and lagged CP/R, so:
//   LEFT  new state = what this very thread computed in its previous iteration         -> kept in registers,
//   DOWN  new state = what lane-1 computed in this iteration                            -> one warp shuffle,
//   RIGHT / UP old state                                                                -> plain loads of the *_in planes.
// Lane 0 of every warp only recomputes the row below the strip (halo), and the sweep starts one column early (halo
// column), so no thread ever waits for another warp: no shared memory, no barrier, redundancy 1/31 + 1/TI.
// The kernel reads the *_in planes and writes the *_out planes (ping-pong).
#pragma once
#include <type_traits>
#include "hf2d_turb.cuh"

struct NbState { double rho, y4, y5, y6, s7, s8, U, V, Tg; };

__device__ __forceinline__ void nb_old(const Params& P, long long m, bool turb, NbState& s) {
    const long long N = P.N;
    s.rho = P.S_in[m]; s.y4 = P.S_in[4 * N + m]; s.y5 = P.S_in[5 * N + m]; s.y6 = P.S_in[6 * N + m];
    s.s7 = turb ? P.S_in[7 * N + m] : 0.; s.s8 = turb ? P.S_in[8 * N + m] : 0.;
    s.U = P.U_in[m]; s.V = P.V_in[m]; s.Tg = P.Tg_in[m];
}
__device__ __forceinline__ double shfl_up1(double v) { return __shfl_up_sync(0xffffffffu, v, 1); }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}
// max / min over the warp of the bit patterns of non-negative doubles (order preserving): two 32-bit REDUX steps
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
    const unsigned hi = (unsigned)(v >> 32), mh = __reduce_max_sync(0xffffffffu, hi);
    const unsigned lo = (hi == mh) ? (unsigned)v : 0u, ml = __reduce_max_sync(0xffffffffu, lo);
    return ((unsigned long long)mh << 32) | ml;
}
__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long v) {
    const unsigned hi = (unsigned)(v >> 32), mh = __reduce_min_sync(0xffffffffu, hi);
    const unsigned lo = (hi == mh) ? (unsigned)v : 0xffffffffu, ml = __reduce_min_sync(0xffffffffu, lo);
    return ((unsigned long long)mh << 32) | ml;
}

// ---- column staging for plain tiles: cp.async (LDGSTS) copies of the warp's segment of every input plane into shared
// memory, issued one column ahead of the sweep, so that no global-load latency is left on the critical path ----
__device__ __forceinline__ void cp_async8(double* smem_dst, const double* gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
#define S2_TILE_DOUBLES (PL_COUNT * TILE_ROWS)          // one staged column of one warp
#ifndef S2_WARPS
#define S2_WARPS 4               // warps (strips) per block
#endif
#ifndef S2_STAGE
#define S2_STAGE 1               // 1: plain tiles stage their columns in shared memory with cp.async
#endif
#define S2_DYN_SMEM_PLAIN (S2_STAGE ? S2_WARPS * 2 * S2_TILE_DOUBLES * 8 : 0)
#define S2_THREADS (32 * S2_WARPS)
// STRIP_ROWS (31, hf2d_stage1.cuh): rows a warp owns; lane 0 is the halo row below them

// tile t covers columns [i_lo + tx*TI, +TI) x rows [sb*STRIP_ROWS, +STRIP_ROWS), t = tx * n_sb + sb; one warp per tile,
// S2_WARPS consecutive entries of the tile list per block
template <class KC>
#ifndef S2_MINB_PLAIN
#define S2_MINB_PLAIN 2     // 255 registers, 8 warps/SM: beats 3 blocks (168 registers, 288 B/thread of spills that miss L1) by 4% (profiles/NOTES.md)
#endif
#ifndef S2_MINB_GEN
#define S2_MINB_GEN 3
#endif
#ifdef S2_MAXNREG      /* tuning builds: explicit register cap instead of a resident-block target */
__global__ void __maxnreg__(S2_MAXNREG)
#else
__global__ void __launch_bounds__(S2_THREADS, KC::plain ? S2_MINB_PLAIN : S2_MINB_GEN)
#endif
k_stage2(Params P, const int* __restrict__ tiles, int n_tiles, int n_sb, int TI, int want_rms) {
    __shared__ double sh_acc[S2_WARPS][2 * NEQ];               // RMS, sumDiv partial sums of the warp (want_rms steps)
    __shared__ unsigned long long sh_u[S2_WARPS][2 * NEQ + 2]; // counts, max DD, dt-min, first error
    __shared__ SmemTables sh_tab;
    if (S2_SMEM_TABLES) smem_tables_fill(sh_tab, P.tab);
    StepState st = *P.st;         // snapshot (the accumulators are only touched through atomics)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t_idx = blockIdx.x * S2_WARPS + warp;
    const int tile = t_idx < n_tiles ? tiles[t_idx] : 0;
    const int tx = tile / n_sb, sb = tile - tx * n_sb;
    const int j = sb * STRIP_ROWS + lane - 1;
    const int i0 = P.i_lo + tx * TI, i1 = t_idx < n_tiles ? min(i0 + TI, P.i_hi) : i0;
    const bool valid_j = j >= 0 && j < P.ny;
    const bool owner_j = lane >= 1 && j < P.ny;
    const long long ny = P.ny, N = P.N;
    if (threadIdx.x < S2_WARPS * 2 * NEQ) (&sh_acc[0][0])[threadIdx.x] = 0.;
    if (lane < 2 * NEQ) sh_u[warp][lane] = 0ull;
    __syncthreads();
    unsigned long long r_dt = ~0ull, r_err = ~0ull;
    NbState carry;   // new state of the left neighbour (previous iteration of this thread)
    carry.rho = carry.y4 = carry.y5 = carry.y6 = carry.s7 = carry.s8 = carry.U = carry.V = carry.Tg = 0.;
    const double yj = (KC::axi && valid_j) ? P.yrow[j] : 1.0;

    // staged columns (plain tiles): two buffers per warp; lane l holds row j of every plane at [plane*TILE_ROWS + l]
    extern __shared__ double dyn_smem[];
    double* const tile0 = dyn_smem + (size_t)warp * 2 * S2_TILE_DOUBLES;
    const int i_first = (i0 > 0 ? i0 - 1 : i0);
    auto stage_column = [&](int col, int buf) {
        double* d = tile0 + buf * S2_TILE_DOUBLES + lane;
        const long long m = (long long)col * ny + j;
        if (valid_j) {
            constexpr int NK = KC::ns ? NEQ : 7;
#pragma unroll
            for (int k = 0; k < NK; k++) {
                cp_async8(d + (PL_S + k) * TILE_ROWS, P.S_in + (long long)k * N + m);
                cp_async8(d + (PL_SN + k) * TILE_ROWS, P.SN + (long long)k * N + m);
            }
            cp_async8(d + PL_U * TILE_ROWS, P.U_in + m); cp_async8(d + PL_V * TILE_ROWS, P.V_in + m); cp_async8(d + PL_TG * TILE_ROWS, P.Tg_in + m);
            cp_async8(d + PL_CP * TILE_ROWS, P.CP_in + m); cp_async8(d + PL_R * TILE_ROWS, P.R_in + m);
            cp_async8(d + PL_KG * TILE_ROWS, P.kg_in + m); cp_async8(d + PL_P * TILE_ROWS, P.p + m);
            if (KC::ns) {
                cp_async8(d + PL_MU * TILE_ROWS, P.mu_in + m); cp_async8(d + PL_MUT * TILE_ROWS, P.mu_t + m);
                cp_async8(d + PL_LAM * TILE_ROWS, P.lam + m); cp_async8(d + PL_LMIN * TILE_ROWS, P.l_min + m);
            }
            if (lane == 31 && j + 1 < P.ny) {   // the row above the strip: up neighbour of the last owner (old state only)
                cp_async8(d + 1 + (PL_S + 0) * TILE_ROWS, P.S_in + m + 1);
#pragma unroll
                for (int k = 4; k < NK; k++) cp_async8(d + 1 + (PL_S + k) * TILE_ROWS, P.S_in + (long long)k * N + m + 1);
                cp_async8(d + 1 + PL_U * TILE_ROWS, P.U_in + m + 1); cp_async8(d + 1 + PL_V * TILE_ROWS, P.V_in + m + 1);
                cp_async8(d + 1 + PL_TG * TILE_ROWS, P.Tg_in + m + 1);
            }
        }
        cp_async_commit();
    };
    int cur = 0;
    constexpr bool STAGED = KC::plain && S2_STAGE;
    if constexpr (STAGED) { if (i_first < i1) { stage_column(i_first, 0); cp_async_wait_all(); __syncwarp(); } }

    // A phase is run with the branch-free DivFast policy on plain tiles and re-run exactly when any lane asks for it
    auto run = [&](auto&& ph) {
        if constexpr (KC::plain) {
            DivFast f; ph(f);
            if (__any_sync(0xffffffffu, f.is_bad())) { DivExact e; ph(e); }
        } else { DivExact e; ph(e); }
    };
    constexpr int NE_C = KC::plain ? (!KC::ns ? 7 : (KC::ttc & TCT_Prandtl_Model) ? 7 : (KC::ttc & TCT_k_eps_Model) ? 9 : (KC::ttc & TCT_SA_Model) ? 8 : 7) : 0;
    constexpr bool TURB_C = KC::plain && KC::ns && (KC::ttc & (TCT_k_eps_Model | TCT_SA_Model)) != 0;

    for (int i = i_first; i < i1; i++) {
        const bool own_col = i >= i0;
        const long long n = valid_j ? (long long)i * ny + j : 0;
        using LD = std::conditional_t<STAGED, TileLoad, GlobalLoad>;
        const double* const tcur = tile0 + cur * S2_TILE_DOUBLES + lane;
        const double* const tnxt = tile0 + (cur ^ 1) * S2_TILE_DOUBLES + lane;
        LD ld = [&]() { if constexpr (STAGED) return TileLoad{tcur}; else return GlobalLoad{P, n}; }();
        if constexpr (STAGED) { if (i + 1 < P.nx) stage_column(i + 1, cur ^ 1); }   // next column: own inputs then, right neighbour now
        uint32_t ct, tt;
        if (KC::plain) { ct = CT_NODE_IS_SET; tt = (uint32_t)KC::ttc | TT_IDXL | TT_IDXR | TT_IDYU | TT_IDYD; }
        else { ct = 0; tt = 0; if (valid_j) { ct = P.ct[n]; tt = P.tt[n]; } }
        const bool act = valid_j && node_active(ct);
        const bool fc = valid_j && !act && is_set(ct, NT_FC) && !is_set(ct, CT_SOLID);
        const bool upd = own_col && owner_j && (act || fc);       // this thread owns the node's Stage 2
        const int ne = num_eq(tt);
        const bool turb2 = KC::ns && (is_set(tt, TCT_k_eps_Model) || is_set(tt, TCT_SA_Model));
        const bool hasturb = KC::ns && (tt & (TCT_k_eps_Model | TCT_SA_Model));
        Local L;
        NbState mine;
        mine.rho = mine.y4 = mine.y5 = mine.y6 = mine.s7 = mine.s8 = mine.U = mine.V = mine.Tg = 0.;
        // ---- commit (:1156-1164) fused with the residual and the locally adaptive blending factor (:1062-1154) ----
        if constexpr (KC::plain) {
            // every equation k < NE_C is advanced and committed; inputs come from the staged column at the point of use
            double bo[NEQ];
#pragma unroll
            for (int k = 0; k < NEQ; k++) { bo[k] = 0.; if (valid_j && k < NE_C && (P.bFF & 1)) bo[k] = P.beta[(long long)k * N + n]; }
            auto phaseA = [&](auto& dv, auto rms) {
                double beta_min = OMIN(P.beta0, st.beta_scn);
                const double bb = beta_min * beta_min;
                // absent species (S[4..6] == 0 on every node of the warp's column: single-gas runs) take no residual and leave
                // beta alone (res == false below): their 6 divisions + 3 square roots are skipped as ONE block, so that the
                // other equations stay a single straight-line stretch
                bool sp_any = false;
#pragma unroll
                for (int k = 4; k < 7; k++) sp_any |= upd && ld.S(k) != 0.;
                sp_any = __any_sync(0xffffffffu, sp_any);
#pragma unroll
                for (int k = 0; k < NE_C; k++) {
                    if (k >= 4 && k < 7 && !sp_any) continue;
                    const double so_k = valid_j ? ld.S(k) : 0., sn_k = valid_j ? ld.SN(k) : 0.;
                    const bool res = upd && so_k != 0.;
                    const double Tmp = res ? so_k : 1.0;
                    const double absDD = sn_k - so_k;
                    const bool big = fabs(Tmp) > 1.e-15;
                    const double q = dv.div(absDD, Tmp, res && big);
                    const double DD = big ? fabs(q) : 1.0;
                    double sqrt_RES = 0;
                    if (P.bFF >= 4) { const double r_ = dv.sqrt(DD, res && big); sqrt_RES = big ? r_ : 0.; }
                    const double xx = (P.bFF >= 4) ? sqrt_RES : (P.bFF >= 2) ? DD * DD : DD;
                    const double cap = (P.bFF & 1) ? (beta_min + bo[k]) * 0.5 : beta_min;
                    const double quo = dv.div(bb, beta_min + xx, res);
                    if (res) P.beta[(long long)k * N + n] = OMIN(cap, quo);
                    if constexpr (decltype(rms)::value) {
                        double c_rms = 0., c_div = 0.;
                        if (res) { if (P.isAlternateRMS) { c_div = Tmp * Tmp; c_rms = absDD; } else { c_rms = DD * DD; } }
                        const double a = warp_sum(c_rms), b = warp_sum(c_div);
                        const unsigned c = __reduce_add_sync(0xffffffffu, res ? 1u : 0u);
                        const unsigned long long d = warp_max_u64(res ? d2bits(DD) : 0ull);
                        if (lane == 0 && c) {
                            sh_acc[warp][k] += a; sh_acc[warp][NEQ + k] += b; sh_u[warp][k] += c;
                            if (d > sh_u[warp][NEQ + k]) sh_u[warp][NEQ + k] = d;
                        }
                    }
                }
            };
            if (want_rms) { DivExact e; phaseA(e, std::true_type()); }     // output steps only: exact policy, fused reduction
            else run([&](auto& dv) { phaseA(dv, std::false_type()); });
#pragma unroll
            for (int k = 0; k < NEQ; k++) L.S[k] = !valid_j ? 0. : (k < NE_C) ? ld.SN(k) : (k < 7 || TURB_C) ? ld.S(k) : 0.;
        } else {
            DivExact dv;
#pragma unroll
            for (int k = 0; k < NEQ; k++) {
                double s = 0., c_rms = 0., c_div = 0.;
                unsigned long long c_dd = 0ull;
                unsigned c_cnt = 0u;
                if (valid_j && (k < 7 || hasturb)) {
                    const long long o = (long long)k * N + n;
                    s = P.S_in[o];
                    if (act && k < ne) {
                        uint32_t c_flag = 0;
                        if (k < 4) c_flag = CT_Rho_CONST << k;
                        else if (k < 4 + NCOMP) c_flag = CT_Y_CONST;
                        else if (turb2) c_flag = TCT_k_CONST << (k - 4 - NCOMP);
                        const bool res = upd && !is_set(ct, c_flag) && s != 0.;   // quirk kept: CT word even for k >= 7 (:1075)
                        const bool com = eq_committed(ct, tt, k, KC::ns);
                        double sn = 0.;
                        if (res || com) sn = P.SN[o];
                        if (res) {
                            const double Tmp = s;
                            const double absDD = sn - s;
                            double DD, sqrt_RES = 0;
                            if (fabs(Tmp) > 1.e-15) { DD = fabs(dv.div(absDD, Tmp)); if (P.bFF >= 4) sqrt_RES = dv.sqrt(DD); } else { DD = 1.0; }
                            double beta_min = OMIN(P.beta0, st.beta_scn);
                            if (is_set(ct, CT_NONREFLECTED)) beta_min = P.nrbc_beta0;
                            double* bp = &P.beta[o];
                            const double bb = beta_min * beta_min;
                            // the six blending-factor functions (deeps2d_core.cpp:1104-1121): min(cap, bb / (beta_min + x))
                            const double xx = (P.bFF >= 4) ? sqrt_RES : (P.bFF >= 2) ? DD * DD : DD;
                            const double cap = (P.bFF & 1) ? (beta_min + *bp) * 0.5 : beta_min;
                            const double quo = dv.div(bb, beta_min + xx);
                            *bp = OMIN(cap, quo);
                            if (P.isAlternateRMS) { c_div = Tmp * Tmp; c_rms = absDD; } else { c_rms = DD * DD; }
                            c_cnt = 1u; c_dd = d2bits(DD);
                        }
                        if (com) s = sn;
                    }
                }
                L.S[k] = s;
                if (want_rms) {   // uniform: residual norms are only read on output steps (deeps2d_core.cpp:1602-1637)
                    const double a = warp_sum(c_rms), b = warp_sum(c_div);
                    const unsigned c = __reduce_add_sync(0xffffffffu, c_cnt);
                    const unsigned long long d = warp_max_u64(c_dd);
                    if (lane == 0 && c) {
                        sh_acc[warp][k] += a; sh_acc[warp][NEQ + k] += b; sh_u[warp][k] += c;
                        if (d > sh_u[warp][NEQ + k]) sh_u[warp][NEQ + k] = d;
                    }
                }
            }
        }
        // ---- the local half of FillNode2D + species chemistry ----
        // (plain tiles may run this twice: local_rest rewrites L.S only idempotently there)
        run([&](auto& dv) {
            if (valid_j) {
                local_rest<KC, true>(dv, ld, P, st, n, ct, tt, L);
                mine.rho = L.S[iRho]; mine.y4 = L.y4; mine.y5 = L.y5; mine.y6 = L.y6; mine.s7 = L.S[7]; mine.s8 = L.S[8];
                mine.U = L.U; mine.V = L.V; mine.Tg = L.Tg;
            }
        });
        // down neighbour's new state: lane-1 just computed it
        NbState nD;
        nD.rho = shfl_up1(mine.rho); nD.y4 = shfl_up1(mine.y4); nD.y5 = shfl_up1(mine.y5); nD.y6 = shfl_up1(mine.y6);
        nD.U = shfl_up1(mine.U); nD.V = shfl_up1(mine.V); nD.Tg = shfl_up1(mine.Tg);
        if (KC::plain ? TURB_C : (KC::ns != 0)) { nD.s7 = shfl_up1(mine.s7); nD.s8 = shfl_up1(mine.s8); }
        else { nD.s7 = 0.; nD.s8 = 0.; }
        // values Stage 2 both reads and overwrites in place are fetched before the phase so that it can be re-run
        double mu_t0 = 0., lam0 = 0.;
        if (KC::ns && upd) { mu_t0 = ld.mu_t(); lam0 = ld.lam(); }
        if constexpr (STAGED) { cp_async_wait_all(); __syncwarp(); }   // column i+1 (issued before phase A) has landed
        const unsigned long long r_dt0 = r_dt, r_err0 = r_err;
        run([&](auto& dv) {
          r_dt = r_dt0; r_err = r_err0;
          if (upd) {
            StepState& stn = st;
            const int keep_init = st.is_init, keep_mu = st.is_mu_t;
            if (fc) { stn.is_init = 0; stn.is_mu_t = 1; }   // NT_FC nodes: FillNode2D(1,0,...) regardless of TurbStartIter (:1331)
            const int n1 = (tt & TT_IDXL) ? 1 : 0, n2 = (tt & TT_IDXR) ? 1 : 0, n3 = (tt & TT_IDYU) ? 1 : 0, n4 = (tt & TT_IDYD) ? 1 : 0;
            SpecNode* sp = (tt & TT_SPEC) ? &P.spec[P.spec_idx[n]] : nullptr;
            // ---- gradients (:1169-1238): left/down neighbours already updated this sweep, right/up not yet ----
            Grads g;
            if (KC::ns) {
                if (act) {
                    NbState own, nR, nL, nU;   // `own`: a missing neighbour aliases the node itself (committed S, old U,V,Tg)
                    own.rho = L.S[iRho]; own.y4 = L.S[4]; own.y5 = L.S[5]; own.y6 = L.S[6]; own.s7 = L.s7c; own.s8 = L.s8c;
                    own.U = L.Uo; own.V = L.Vo; own.Tg = L.Tgo;
                    if constexpr (STAGED) {   // old state of the right / up neighbours from the staged columns
                        nR.rho = tnxt[PL_S * TILE_ROWS]; nR.y4 = tnxt[(PL_S + 4) * TILE_ROWS]; nR.y5 = tnxt[(PL_S + 5) * TILE_ROWS]; nR.y6 = tnxt[(PL_S + 6) * TILE_ROWS];
                        nR.s7 = TURB_C ? tnxt[(PL_S + 7) * TILE_ROWS] : 0.; nR.s8 = TURB_C ? tnxt[(PL_S + 8) * TILE_ROWS] : 0.;
                        nR.U = tnxt[PL_U * TILE_ROWS]; nR.V = tnxt[PL_V * TILE_ROWS]; nR.Tg = tnxt[PL_TG * TILE_ROWS];
                        nU.rho = tcur[1 + PL_S * TILE_ROWS]; nU.y4 = tcur[1 + (PL_S + 4) * TILE_ROWS]; nU.y5 = tcur[1 + (PL_S + 5) * TILE_ROWS]; nU.y6 = tcur[1 + (PL_S + 6) * TILE_ROWS];
                        nU.s7 = TURB_C ? tcur[1 + (PL_S + 7) * TILE_ROWS] : 0.; nU.s8 = TURB_C ? tcur[1 + (PL_S + 8) * TILE_ROWS] : 0.;
                        nU.U = tcur[1 + PL_U * TILE_ROWS]; nU.V = tcur[1 + PL_V * TILE_ROWS]; nU.Tg = tcur[1 + PL_TG * TILE_ROWS];
                    } else {
                        if (n2) nb_old(P, n + ny, hasturb, nR); else nR = own;
                        if (n3) nb_old(P, n + 1, hasturb, nU); else nU = own;
                    }
                    if (n1) nL = carry; else nL = own;
                    if (!n4) nD = own;
                    const double n_n = OMAX(n1 + n2, 1), m_m = OMAX(n3 + n4, 1);
                    const double dxn = (n1 + n2 == 2) ? P.dxn[1] : P.dxn[0], dym = (n3 + n4 == 2) ? P.dyn[1] : P.dyn[0];   // dx_1/n_n, dy_1/m_m (:1169)
                    double aR = nR.rho, aL = nL.rho, aU = nU.rho, aD = nD.rho;
                    const bool gx = !is_set(ct, CT_dYdx_NULL), gy = !is_set(ct, CT_dYdy_NULL);
                    // droYdx/droYdy keep their stored value when the species Neumann flag is set; nothing reads the
                    // species entries in that case except FillNode2D below -> they are kept in the special-node table
                    if (sp) { for (int q = 0; q < 4; q++) { g.droYdx[q] = sp->grad[q]; g.droYdy[q] = sp->grad[4 + q]; } }
                    else { for (int q = 0; q < 4; q++) { g.droYdx[q] = 0.; g.droYdy[q] = 0.; } }
                    g.droYdx[3] = 0.; g.droYdy[3] = 0.;
                    if (gx) { g.droYdx[0] = (nR.y4 - nL.y4) * dxn; aR -= nR.y4; aL -= nL.y4; }
                    if (gy) { g.droYdy[0] = (nU.y4 - nD.y4) * dym; aU -= nU.y4; aD -= nD.y4; }
                    if (gx) { g.droYdx[1] = (nR.y5 - nL.y5) * dxn; aR -= nR.y5; aL -= nL.y5; }
                    if (gy) { g.droYdy[1] = (nU.y5 - nD.y5) * dym; aU -= nU.y5; aD -= nD.y5; }
                    if (gx) { g.droYdx[2] = (nR.y6 - nL.y6) * dxn; aR -= nR.y6; aL -= nL.y6; }
                    if (gy) { g.droYdy[2] = (nU.y6 - nD.y6) * dym; aU -= nU.y6; aD -= nD.y6; }
                    if (gx) g.droYdx[3] = (aR - aL) * dxn;
                    if (gy) g.droYdy[3] = (aU - aD) * dym;
                    g.dkdx = g.dkdy = g.depsdx = g.depsdy = 0.;
                    if (sp) { g.dkdx = sp->grad[14]; g.dkdy = sp->grad[15]; g.depsdx = sp->grad[16]; g.depsdy = sp->grad[17]; }
                    if (is_set(ct, CT_WALL_NO_SLIP) || is_set(ct, CT_WALL_LAW)) {
                        g.dUdx = (nR.U * n1 - nL.U * n2) * dxn; g.dVdx = (nR.V * n1 - nL.V * n2) * dxn;
                        g.dUdy = (nU.U * n3 - nD.U * n4) * dym; g.dVdy = (nU.V * n3 - nD.V * n4) * dym;
                        if (is_set(tt, TCT_k_eps_Model)) {
                            g.dkdx = dv.div((nR.s7 * n1 - nL.s7 * n2) * dxn, L.S[iRho]); g.depsdx = dv.div((nR.s8 * n1 - nL.s8 * n2) * dxn, L.S[iRho]);
                            g.dkdy = dv.div((nU.s7 * n3 - nD.s7 * n4) * dym, L.S[iRho]); g.depsdy = dv.div((nU.s8 * n3 - nD.s8 * n4) * dym, L.S[iRho]);
                        } else if (is_set(tt, TCT_SA_Model)) {
                            g.dkdx = dv.div((nR.s7 * n1 - nL.s7 * n2) * dxn, L.S[iRho]); g.dkdy = dv.div((nU.s7 * n3 - nD.s7 * n4) * dym, L.S[iRho]);
                        }
                    } else {
                        g.dUdx = (nR.U - nL.U) * dxn; g.dVdx = (nR.V - nL.V) * dxn;
                        g.dUdy = (nU.U - nD.U) * dym; g.dVdy = (nU.V - nD.V) * dym;
                        if (is_set(tt, TCT_k_eps_Model)) {
                            g.dkdx = dv.div((nR.s7 - nL.s7) * dxn, L.S[iRho]); g.depsdx = dv.div((nR.s8 - nL.s8) * dxn, L.S[iRho]);
                            g.dkdy = dv.div((nU.s7 - nD.s7) * dym, L.S[iRho]); g.depsdy = dv.div((nU.s8 - nD.s8) * dym, L.S[iRho]);
                        } else if (is_set(tt, TCT_SA_Model)) {
                            g.dkdx = dv.div((nR.s7 - nL.s7) * dxn, L.S[iRho]); g.dkdy = dv.div((nU.s7 - nD.s7) * dym, L.S[iRho]);
                        }
                    }
                    g.dTdx = (nR.Tg - nL.Tg) * dxn; g.dTdy = (nU.Tg - nD.Tg) * dym;
                    // the only gradients anything reads after the step: dUdy, dVdx of the wall nodes (Recalc_y_plus when a
                    // run is restarted from the swap file, deeps2d_core.cpp:2364-2386) -> kept in the special-node table for
                    // hf2d_gpu_download_aos; these two slots are read back by nobody on an active node
                    if constexpr (!KC::plain) {
                        if (sp && (is_set(ct, CT_WALL_NO_SLIP) || is_set(ct, CT_WALL_LAW))) { sp->grad[9] = g.dUdy; sp->grad[10] = g.dVdx; }
                    }
                } else {   // NT_FC node: FillNode2D runs on the gradients stored in the record (never refreshed)
                    for (int q = 0; q < 4; q++) { g.droYdx[q] = sp->grad[q]; g.droYdy[q] = sp->grad[4 + q]; }
                    g.dUdx = sp->grad[8]; g.dUdy = sp->grad[9]; g.dVdx = sp->grad[10]; g.dVdy = sp->grad[11];
                    g.dTdx = sp->grad[12]; g.dTdy = sp->grad[13];
                    g.dkdx = sp->grad[14]; g.dkdy = sp->grad[15]; g.depsdx = sp->grad[16]; g.depsdy = sp->grad[17];
                }
            }
#include "hf2d_stage2_fill.inc"
            stn.is_init = keep_init; stn.is_mu_t = keep_mu;
          }
        });
        carry = mine;
        if constexpr (STAGED) { __syncwarp(); cur ^= 1; }   // every lane is done with this column's buffer before it is refilled
    }
#include "hf2d_stage2_reduce.inc"
}
