// hf2d_heat.cuh -- CalcHeatOnWallSources (deeps2d_core.cpp:2679-2833) as a sparse kernel over the wall-node list.
//
// The reference sweeps i-outer/j-inner; every wall node adjacent to a SOLID node deposits a conductive flux on that
// solid node's Q_conv (averaging with what earlier wall nodes of the sweep left there) and takes
// SrcAdd[RhoE] = -dt*Q_conv/d from the running value.  A solid node (si,sj) is visited, in sweep order, by
// (si-1,sj) [as its Right], (si,sj-1) [Up], (si,sj+1) [Down], (si+1,sj) [Left]; each wall thread replays the (at
// most 3) visits that precede its own, so the result is order-exact without any serialisation.
// FlowNode2D::UpNode..RightNode are never assigned (hyper_flow_node.hpp:154-157) => num_near_nodes == 1 always.
#pragma once
#include "hf2d_dev.cuh"

// ghost_ok: the ghost columns of an X-slab hold the neighbour slabs' NEW Tg, lam and lam_t (the phase-0 exchange has run:
// hf2d_gpu_step_part S2_FINISH), so a wall node beyond the seam visits a solid node exactly as it does in the undivided
// grid.  (Not covered: a solid node IN a ghost column whose other-side visitor lies two columns beyond the seam -- a solid
// feature one cell thick straddling a seam.)
__device__ __forceinline__ bool heat_visitor(const Params& P, int vi, int vj, long long& v, int ghost_ok) {
    if (vi < (ghost_ok ? 0 : P.i_lo) || vi >= (ghost_ok ? P.nx : P.i_hi) || vj < 0 || vj >= P.ny) return false;
    v = (long long)vi * P.ny + vj;
    const uint32_t c = P.ct[v];
    if (is_set(c, CT_SOLID)) return false;
    return is_set(c, CT_WALL_LAW) || is_set(c, CT_WALL_NO_SLIP);
}

// running Q_conv of solid node (si,sj) after the visits up to and including visitor number `upto` (0..3)
__device__ __forceinline__ double heat_chain(const Params& P, const double* __restrict__ Tg, const double* __restrict__ lam, int si, int sj, int upto, int ghost_ok) {
    const long long s = (long long)si * P.ny + sj;
    const double Ts = Tg[s];
    const int vi[4] = {si - 1, si, si, si + 1}, vj[4] = {sj, sj - 1, sj + 1, sj};
    double Q = 0.;
    for (int q = 0; q <= upto; q++) {
        long long v;
        if (!heat_visitor(P, vi[q], vj[q], v, ghost_ok)) continue;
        const double d = (q == 0 || q == 3) ? P.dx : P.dy;
        double lam_eff = lam[v] + P.spec[P.spec_idx[v]].lam_t;
        lam_eff = lam_eff / 1;
        if (Q > 0.) Q = (Q - lam_eff * (Ts - Tg[v]) / d) * 0.5;
        else Q = -lam_eff * (Ts - Tg[v]) / d;
    }
    return Q;
}

// Tg, lam: the planes K1 just wrote (the heat source runs after the sweep, deeps2d_core.cpp:1402)
__global__ void k_wall_heat(Params P, const double* __restrict__ Tg, const double* __restrict__ lam, const int* __restrict__ wall_list, int n_wall, int ghost_ok) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_wall) return;
    SpecNode& sp = P.spec[wall_list[t]];
    const long long n = sp.node;
    const int i = (int)(n / P.ny), j = (int)(n % P.ny);
    if (i < P.i_lo || i >= P.i_hi) return;
    const double dt = P.st->dt;
    // direction order of the reference: Down, Up, Left, Right; visitor slot of THIS node on that solid node
    const int si[4] = {i, i, i - 1, i + 1}, sj[4] = {j - 1, j + 1, j, j};
    const int slot[4] = {2, 1, 3, 0};
    for (int d = 0; d < 4; d++) {
        sp.Q_conv[d] = 0.;
        if (si[d] < 0 || si[d] >= P.nx || sj[d] < 0 || sj[d] >= P.ny) continue;
        const long long s = (long long)si[d] * P.ny + sj[d];
        if (!is_set(P.ct[s], CT_SOLID)) continue;
        const double Q = heat_chain(P, Tg, lam, si[d], sj[d], slot[d], ghost_ok);
        sp.Q_conv[d] = Q;
        sp.SrcAdd[iRhoE] = -dt * Q / ((d < 2) ? P.dy : P.dx);
    }
}
