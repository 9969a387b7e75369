/* TEST INFRASTRUCTURE ONLY -- CPU oracle for the DEEPS2D explicit time step.
 *
 * Plain-C, single-threaded restatement of the reference's serial solver sweep
 * (OpenHyperFLOW2D/libDEEPS2D/deeps2d_core.cpp:735-1642, node physics from
 * libOpenHyperFLOW2D/hyper_flow_node.hpp:374-957).  It is the checker the CUDA
 * path is compared with; it is NEVER linked into, imported by or called from
 * the product library (openhyperflow2d_b200/).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may use it.
 *
 * Parity status: PINNED.  tests/test_oracle_vs_reference.py compares every field of
 * this oracle with the raw AoS swap files written by the unmodified reference
 * binary (oracle/_ref/ohf_dump, built by oracle/Makefile) -- committed as
 * tests/golden/*.npz together with the generating script tests/golden/make_golden.py.
 */
#ifndef DEEPS2D_ORACLE_H
#define DEEPS2D_ORACLE_H
#include <stdint.h>

#define OHF_NEQ   9   /* FlowNode2D::NumEq = 6 + NUM_COMPONENTS, hyper_flow_node.hpp:255 */
#define OHF_NCOMP 3   /* NUM_COMPONENTS, hyper_flow_node.hpp:33 */

/* One grid node.  Field names follow FlowNode2D (hyper_flow_node.hpp:138-191) and
 * FlowNodeTurbulence2D (hyper_flow_turbulence.hpp:107-133); the memory layout is
 * the oracle's own (all 8-byte members), conversion from/to the reference's
 * 1248-byte record is done by name in python (openhyperflow2d_b200/aos.py). */
typedef struct ohf_node {
    double   S[OHF_NEQ], dSdx[OHF_NEQ], dSdy[OHF_NEQ];
    double   SN[OHF_NEQ];                 /* FlowNodeCore2D shadow ("NextNode->S") */
    uint64_t TurbType;
    double   l_min, y_plus, Re_local, mu_t, lam_t, dkdx, dkdy, depsdx, depsdy;
    double   y;
    double   p;
    int64_t  idXl, idYu, idXr, idYd;
    uint64_t CT;
    double   beta[OHF_NEQ];
    double   Q_conv, time, k, R, lam, mu, CP, Diff, Tf;
    double   A[OHF_NEQ], B[OHF_NEQ], F[OHF_NEQ], RX[OHF_NEQ], RY[OHF_NEQ], Src[OHF_NEQ], SrcAdd[OHF_NEQ];
    double   Tg, U, V, Y[OHF_NCOMP + 1];
    double   Uw, Vw;
    double   droYdx[OHF_NCOMP + 1], droYdy[OHF_NCOMP + 1];
    double   dUdx, dUdy, dVdx, dVdy, dTdx, dTdy;
    double   BGX, BGY;
} ohf_node;

typedef struct ohf_table { int n; const double* x; const double* y; } ohf_table;

/* property-table slots, order fixed: species index s = 0 fuel, 1 oxidiser, 2 products, 3 air */
enum { OHF_T_CP = 0, OHF_T_LAM = 4, OHF_T_MU = 8, OHF_T_CFL = 12, OHF_T_BETA = 13, OHF_NTABLES = 14 };

typedef struct ohf_cfg {
    int    nx, ny;
    double dx, dy;
    double CFL, beta0, nrbc_beta0;
    int    bFF;               /* BlendingFactorFunction, deeps2d_core.hpp:66 */
    int    problem_type;      /* 0 Euler, 1 NS  (SolverMode) */
    int    flow_type;         /* 0 flat, 1 axisymmetric (FlowNode2D::FT) */
    double Hu[4];
    int    TurbStartIter, TurbExtModel, isTurbulenceReset;
    double SigW, SigF, delta_bl, turb_intensity;
    int    isAdiabaticWall, isAlternateRMS;
    unsigned AddSrcStartIter;
    int    MonitorIndex;
    double ExitMonitorValue;
    double K0;
    double Rgas[4];           /* R_Fuel, R_OX, R_cp, R_air */
    ohf_table tab[OHF_NTABLES];
} ohf_cfg;

typedef struct ohf_step_report {
    double dt;                /* dt used by this step */
    double dt_min;            /* running minimum after this step (serial rule, deeps2d_core.cpp:803,1326) */
    double sum_RMS[OHF_NEQ];  /* per-equation residual norm, deeps2d_core.cpp:1539-1574 */
    double max_RMS;
    int    k_max_RMS;
    double DD_max[OHF_NEQ];
    int    i_c, j_c;
    int    err;               /* 1 if Tg<0 was hit (deeps2d_core.cpp:1246); err_i, err_j = node */
    int    err_i, err_j;
} ohf_step_report;

#ifdef __cplusplus
extern "C" {
#endif
/* sizeof(ohf_node) and offsets, so python can build a matching dtype without guessing */
int    ohf_node_bytes(void);
double ohf_table_getval(const ohf_table* t, double x);
/* One explicit step over the whole grid (single sub-domain, serial semantics).
 * `global_iter` = iter+last_iter of the reference; `is_src_add` = FlowNode2D::isSrcAdd for the
 * current cycle; `dt_min` in/out running minimum (dt_min[0] of the reference).            */
int ohf_step(ohf_node* J, const ohf_cfg* cfg, unsigned global_iter, int is_src_add,
             double global_time, double* dt_min, ohf_step_report* rep);
/* n steps of one cycle: iter = 0..n-1 with last_iter fixed (deeps2d_core.cpp:723-1642). */
int ohf_run_cycle(ohf_node* J, const ohf_cfg* cfg, unsigned last_iter, int nsteps,
                  double global_time, double* dt_min, ohf_step_report* rep_last, double* time_part);
#ifdef __cplusplus
}
#endif
/* Wall nodes and wall distance (GetWallNodes, deeps2d_core.cpp:2081-2104; SetMinDistanceToWall2D, :4783-4832):
 * planar, x-major.  Pinned by tests/test_zzz_wall_distance.py against l_min / i_wall / j_wall of the reference's own
 * pre-step images (tests/golden: the reference runs both at start-up of every Navier-Stokes case). */
long ohf_get_wall_nodes(const uint64_t* ct, int nx, int ny, int* out_ij);
void ohf_set_min_distance_to_wall(uint64_t* ct, const double* Tg, const double* p, int nx, int ny, double dx, double dy,
                                  double x0, const int* wall_ij, long n_wall, double* l_min, int* i_wall, int* j_wall,
                                  unsigned char* status);

/* Recalc_y_plus (deeps2d_core.cpp:2364-2386), planar; pinned by tests/test_zzz_wall_distance.py against the y_plus member of
 * the reference's start-up images of RESTARTED runs (tests/golden/restart: the only ones with non-zero wall gradients). */
void ohf_recalc_y_plus(const uint64_t* ct, int nx, int ny, double dx, double dy, const int* i_wall, const int* j_wall,
                       const double* dUdy, const double* dVdx, const double* mu, const double* rho, double* y_plus);

#endif
