/* hf2d_gpu.h -- C ABI of the B200-native DEEPS2D time step (libhf2d_gpu.so).
 *
 * Drop-in boundary for ONE path of OpenHyperFLOW2D: the inner `do { ... } while(iter < Nstep)` loop of
 * `DEEPS2D_Run` (libDEEPS2D/deeps2d_core.cpp:735-1642; declaration deeps2d_core.hpp:206-212) -- Stage 1,
 * Stage 2 (FlowNode2D::FillNode2D / TurbModRANS2D, hyper_flow_node.hpp:374-957), CalcChemicalReactions
 * (deeps2d_core.cpp:4697-4780), CalcHeatOnWallSources (:2679-2833), the CFL dt reduction (:1318-1326) and the
 * residual reduction (:1089-1153, :1539-1574).  The reference has no FFI: the seam is that C++ function plus
 * the file-scope globals it reads (deeps2d_core.hpp:118-151).  A reference-side shim therefore
 *   1. fills `hf2d_cfg` from those globals                         (replaces: reading them inside the loop),
 *   2. describes its FlowNode2D record with `hf2d_node_layout`     (offsetof on hyper_flow_node.hpp:138-191),
 *   3. hf2d_gpu_upload_aos(J->GetMatrixPtr())                      (replaces: SubDomainArray views, hf2d_start.cpp:310-322),
 *   4. hf2d_gpu_step(Nstep) once per output cycle                  (replaces: deeps2d_core.cpp:735-1642),
 *   5. hf2d_gpu_download_aos(J->GetMatrixPtr()) before the per-cycle writers run (DataSnapshot :1773, swap-file
 *      sync :1818-1849, libOutCFD post-processing :1790-1813).
 * INTEGRATION.md shows that shim.  Plain C types only; every function returns 0 on success or an HF2D_E* code;
 * hf2d_gpu_last_error() gives a message.  One host thread per handle; one handle per GPU (X-slab).
 * There is no CPU fallback: without a CUDA device every entry point but the layout/size helpers fails.
 */
#ifndef HF2D_GPU_H
#define HF2D_GPU_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HF2D_NEQ      9    /* FlowNode2D::NumEq = 6 + NUM_COMPONENTS (hyper_flow_node.hpp:255) */
#define HF2D_NTABLES 14

enum hf2d_status {
    HF2D_OK = 0,
    HF2D_EINVAL = 1,       /* bad argument / unsupported configuration                                   */
    HF2D_ECUDA = 2,        /* CUDA runtime error (no device, out of memory, launch failure)               */
    HF2D_EUNSTABLE = 3,    /* Tg < 0 met (deeps2d_core.cpp:1246): report.err_i/err_j hold the first node  */
    HF2D_ESTATE = 4        /* call out of order (e.g. step before upload)                                 */
};

/* property/scenario table, Table::GetVal semantics (obj_data/obj_data.cpp:1822-1860); n == 0 is the zero table */
typedef struct hf2d_table { int n; const double* x; const double* y; } hf2d_table;

/* table slots: species s = 0 fuel, 1 oxidiser, 2 combustion products, 3 air (ChemicalReactionsModelData2D,
 * hyper_flow_bound.hpp:44-76); CFL_Scenario / beta_Scenario: deeps2d_core.cpp:737-738 */
enum { HF2D_T_CP = 0, HF2D_T_LAM = 4, HF2D_T_MU = 8, HF2D_T_CFL = 12, HF2D_T_BETA = 13 };

typedef struct hf2d_cfg {
    int    nx, ny;              /* MaxX, MaxY of this slab (ghost columns included when n_ranks > 1)        */
    double dx, dy;              /* FlowNode2D::dx, dy                                                       */
    double dt;                  /* initial dt = dt_min[0] at entry (deeps2d_core.cpp:660)                   */
    double CFL, beta0, nrbc_beta0;
    int    bFF;                 /* BlendingFactorFunction 0..5 (deeps2d_core.hpp:66-72)                     */
    int    problem_type;        /* SolverMode: 0 SM_EULER, 1 SM_NS                                          */
    int    flow_type;           /* FlowType: 0 FT_FLAT, 1 FT_AXISYMMETRIC                                   */
    double Hu[4];               /* FlowNode2D::Hu                                                           */
    int    TurbStartIter, TurbExtModel, isTurbulenceReset;
    double SigW, SigF, delta_bl;
    double turb_intensity;      /* FlowNodeTurbulence2D::I (0.005)                                          */
    int    isAdiabaticWall, isAlternateRMS;
    unsigned AddSrcStartIter;
    unsigned last_iter;         /* global iteration count at entry                                          */
    int    MonitorIndex;
    double ExitMonitorValue;
    double K0;                  /* chemical_reactions.K0                                                    */
    double Rgas[4];             /* R_Fuel, R_OX, R_cp, R_air                                                */
    hf2d_table tab[HF2D_NTABLES];
    int    dt_rule;             /* 0 = serial running minimum (deeps2d_core.cpp:803,1326; the oracle's rule),
                                   1 = per-step minimum (MPI build, :806-809,850)                           */
    int    rank, n_ranks;       /* X-slab index / count; 0,1 for a single GPU                               */
    int    NOutStep;            /* the reference reads the residual norms only on steps with iter % NOutStep == 0
                                   (SaveRMS / progress line, deeps2d_core.cpp:1602-1637) and after the last step of
                                   a cycle (exit monitor, :1870-1880): hf2d_gpu_step() reduces them on exactly those
                                   steps of a call.  0 or 1 = every step.                                       */
} hf2d_cfg;

/* Byte offsets of the members of the caller's FlowNode2D record (offsetof on the reference's own header).
 * int members (idXl..) are 4-byte ints, CT/TurbType 8-byte unsigned, everything else double.             */
typedef struct hf2d_node_layout {
    size_t node_bytes;
    size_t S, dSdx, dSdy, TurbType, l_min, y_plus, mu_t, lam_t, dkdx, dkdy, depsdx, depsdy, y, p,
           idXl, idYu, idXr, idYd, CT, beta, Q_conv, time, k, R, lam, mu, CP, Diff, Tf,
           A, B, F, Src, SrcAdd, Tg, U, V, Y, Uw, Vw, droYdx, droYdy, dUdx, dUdy, dVdx, dVdy, dTdx, dTdy, BGX, BGY;
} hf2d_node_layout;

typedef struct hf2d_step_report {
    double   dt;                 /* dt used by the last step                                               */
    double   dt_min;             /* value the next step will use                                           */
    double   time_part;          /* sum of dt over the steps of this call (CurrentTimePart, :1602)         */
    double   sum_RMS[HF2D_NEQ];  /* residual norms of the last step (:1539-1574)                           */
    double   max_RMS;
    int      k_max_RMS;
    double   DD_max[HF2D_NEQ];
    unsigned iter;               /* global iteration count after the call                                  */
    int      err, err_i, err_j;  /* Tg<0: first offending node in sweep order                              */
    double   gpu_ms;             /* device time of the step loop (CUDA events)                             */
    uint64_t kernel_launches;    /* kernels launched by this call                                          */
} hf2d_step_report;

typedef struct hf2d_gpu hf2d_gpu;   /* opaque handle */

const char* hf2d_gpu_last_error(void);
int  hf2d_gpu_device_count(void);

/* Create a solver for one slab on CUDA device `device`.  Tables are copied. */
int  hf2d_gpu_create(const hf2d_cfg* cfg, int device, hf2d_gpu** out);
int  hf2d_gpu_destroy(hf2d_gpu* h);

/* Host AoS grid image J[x*ny+y] -> SoA arrays in HBM (also (re)builds the boundary-node lists and zeroes the
 * FlowNodeCore2D shadow exactly as hf2d_start.cpp:318 does).  `pinned` != 0 promises page-locked memory. */
int  hf2d_gpu_upload_aos(hf2d_gpu* h, const void* J, const hf2d_node_layout* layout, int pinned);
/* SoA arrays -> host AoS image; only members the step owns are written, everything else in J is left alone. */
int  hf2d_gpu_download_aos(hf2d_gpu* h, void* J, const hf2d_node_layout* layout, int pinned);
/* Page-lock / release the caller's grid image (cudaHostRegister) so that the upload and the per-cycle downloads run
 * at PCIe speed; worth it when J lives through many output cycles (the reference allocates it once, hf2d_start.cpp:60-146).
 * Optional: pageable memory works, slower. */
int  hf2d_gpu_pin_host(void* J, size_t bytes);
int  hf2d_gpu_unpin_host(void* J);

/* `nsteps` explicit time steps (one output cycle of the reference: iter = 0..nsteps-1 with last_iter fixed,
 * isSrcAdd evaluated once at cycle start, deeps2d_core.cpp:727-731).  Asynchronous on the handle's stream
 * until the report is read at the end.  `want_rms` = 0 skips the residual reduction read-back. */
int  hf2d_gpu_step(hf2d_gpu* h, int nsteps, hf2d_step_report* report);
/* further steps of the SAME cycle (isSrcAdd and time_part carry on): lets a shim report every NOutStep steps (:1602-1637) */
int  hf2d_gpu_step_continue(hf2d_gpu* h, int nsteps, hf2d_step_report* report);
/* p and Tg of one node (monitor points, :1576-1600) */
int  hf2d_gpu_read_node(hf2d_gpu* h, int i, int j, double* p, double* Tg);

/* One reference output cycle through host memory: upload + nsteps + download (what a shim that keeps J on the
 * host between cycles does).  Used for the end-to-end metric. */
int  hf2d_gpu_run_cycle(hf2d_gpu* h, void* J, const hf2d_node_layout* layout, int pinned, int nsteps,
                        hf2d_step_report* report);

/* Device-resident access for callers that keep the grid in HBM (benchmarks, multi-GPU drivers).
 * name: "S","SN","beta","A","B","F","Src" (k = equation) or "U","V","Tg","p","k","R","CP","lam","mu","mu_t",
 * "l_min","y_plus" (k = 0).  Returns the CURRENT device pointer of that plane (nx*ny doubles, x-major). */
int  hf2d_gpu_field_ptr(hf2d_gpu* h, const char* name, int k, void** dev_ptr);
int  hf2d_gpu_active_cells(hf2d_gpu* h, long long* n_active, long long* n_total);
/* strip-sweep tiles of the current grid: `plain` tiles (unflagged interior nodes only) run the specialised kernels,
 * `general` tiles the flag-decoding ones; tile = tile_cols columns x tile_rows rows */
int  hf2d_gpu_tile_counts(hf2d_gpu* h, int* n_plain, int* n_general, int* tile_cols, int* tile_rows);
/* algorithmic bytes per active cell per step of the current configuration */
int  hf2d_gpu_bytes_per_cell(hf2d_gpu* h, double* bytes);

/* Multi-GPU (X-slabs, one process per GPU): pack the columns the neighbours need into `buf` / consume received
 * ones.  side: 0 = left neighbour, 1 = right neighbour.  phase: 0 = full edge columns after Stage 2 (both ways),
 * 1 = NEW state (S, U, V, Tg) of the edge column between K1 and K2 (left -> right), 2 = S of the edge column after the
 * Neumann pre-pass (left -> right; makes Neumann chains that cross a seam exact) -- see DESIGN.md "seams".  `buf` is device memory of hf2d_gpu_halo_bytes() bytes. */
size_t hf2d_gpu_halo_bytes(hf2d_gpu* h, int phase);
int  hf2d_gpu_halo_pack(hf2d_gpu* h, int side, int phase, void* dev_buf);
int  hf2d_gpu_halo_unpack(hf2d_gpu* h, int side, int phase, const void* dev_buf);
/* Split step for drivers that exchange halos between phases: begin (dt, scenario), stage1, stage2, end. */
int  hf2d_gpu_step_begin(hf2d_gpu* h);
int  hf2d_gpu_step_prepass(hf2d_gpu* h);   /* optional: the ordered Neumann pre-pass alone; stage1 then skips it (halo phase 2) */
int  hf2d_gpu_step_stage1(hf2d_gpu* h);
int  hf2d_gpu_step_stage2(hf2d_gpu* h);
int  hf2d_gpu_step_end(hf2d_gpu* h, hf2d_step_report* report /* may be NULL: no read-back */);
/* dt of the running minimum as device pointer (one double) so a driver can all-reduce(min) it in place */
/* Overlapped X-slab step: the same step cut into parts that the driver queues on two stream "lanes" (0 = the handle's
 * main stream, 1 = its high-priority side stream, hf2d_gpu_side_stream) so that seam traffic hides behind the bulk:
 *     begin | fork | lane 1: PREPASS_LOCAL [recv left, unpack(0,2)] PREPASS [pack(1,2), send right] S1_GENERAL | lane 0: S1_PLAIN
 *           | join | S1_FINISH | fork
 *           | lane 1: pack(1,1) pack(0,2) <-> unpack(0,1) unpack(1,2)  S2_EDGE  pack(.,0) <-> unpack(.,0) | lane 0: S2_INNER
 *           | join | S2_FINISH | dt all-reduce | end
 * (openhyperflow2d_b200/multigpu.py is that driver over torch.distributed; replaces the once-per-step blocking exchange
 * of the reference's MPI build, deeps2d_core.cpp:1336-1400).  Between S2_EDGE and S2_FINISH halo phase 0 addresses the
 * NEW state.  The first hf2d_gpu_edge_tile_counts() tiles of each class are the ones that read a ghost column. */
enum hf2d_step_part_id {
    HF2D_PART_PREPASS = 1, HF2D_PART_S1_PLAIN = 2, HF2D_PART_S1_GENERAL = 3, HF2D_PART_S1_FINISH = 4,
    HF2D_PART_S2_INNER = 5, HF2D_PART_S2_EDGE = 6, HF2D_PART_S2_FINISH = 7,
    HF2D_PART_AHEAD = 9,          /* on lane 1 after the phase-0 exchange, exact seams: the NEXT step's pre-pass + hand-off run now, hidden
                                     behind S2_INNER (they touch only nodes of general tiles, all of which have finished Stage 2);
                                     the next step then starts at S1_GENERAL.  Not before a download: S would hold the overwrite. */
    HF2D_PART_PREPASS_LOCAL = 8   /* optional, before the phase-2 receive: the part of the pre-pass that does not wait for it;
                                     PREPASS then runs only what depends on the received column */
};
int  hf2d_gpu_step_part(hf2d_gpu* h, int part, int lane, int want_rms);
int  hf2d_gpu_lane_fork(hf2d_gpu* h);    /* lane 1 waits for everything queued on lane 0 so far */
int  hf2d_gpu_lane_join(hf2d_gpu* h);    /* lane 0 waits for everything queued on lane 1 so far */
void* hf2d_gpu_side_stream(hf2d_gpu* h); /* cudaStream_t of lane 1 */
int  hf2d_gpu_halo_pack_lane(hf2d_gpu* h, int side, int phase, void* dev_buf, int lane);
int  hf2d_gpu_halo_unpack_lane(hf2d_gpu* h, int side, int phase, const void* dev_buf, int lane);
int  hf2d_gpu_edge_tile_counts(hf2d_gpu* h, int* n_plain_edge, int* n_general_edge);
int  hf2d_gpu_dt_ptr(hf2d_gpu* h, void** dev_ptr);
void* hf2d_gpu_stream(hf2d_gpu* h);   /* cudaStream_t the handle launches on */
int  hf2d_gpu_set_stream(hf2d_gpu* h, void* cuda_stream);   /* launch on the caller's stream instead */
int  hf2d_gpu_dt_export(hf2d_gpu* h, void* dev_double);     /* running dt minimum -> caller's device double */
int  hf2d_gpu_dt_import(hf2d_gpu* h, const void* dev_double);
/* per-kernel device time of the three plain-tile launches (slot 0 = K1a k_stage1_plain, 1 = K1b k_point, 2 = K2 k_flux),
 * CUDA events on the launching stream; enable != 0 (re)starts the collection (at most 256 launches per slot are kept) */
int  hf2d_gpu_set_timing(hf2d_gpu* h, int enable);
int  hf2d_gpu_get_timing(hf2d_gpu* h, double ms_mean[3], int count[3]);
unsigned long long hf2d_gpu_launch_count(hf2d_gpu* h);      /* kernels launched through this handle so far */

/* GlobalTime of the running output cycle.  The reference stamps it on every swept node (FlowNode2D::time,
 * deeps2d_core.cpp:867) and a run restarted from the swap file takes its GlobalTime from node (0,0) (:4618-4621).
 * Optional: once set, hf2d_gpu_download_aos writes it into the `time` member of the swept nodes; never set, that member
 * comes back as uploaded.  Call it before the download of each cycle with the cycle's GlobalTime (before += CurrentTimePart). */
int  hf2d_gpu_set_global_time(hf2d_gpu* h, double global_time);

/* Gradient members of the records (droYdx[4], droYdy[4], dUdx .. dTdy, dkdx .. depsdy; deeps2d_core.cpp:1169-1238).  They are
 * intermediates of a step -- never stored in HBM -- but a run RESTARTED from the swap file re-runs FillNode2D(0,1) on every
 * node with the values its records hold (InitDEEPS2D, deeps2d_core.cpp:4565; ScanArea with isTurbulenceReset, :2186-2205).
 * enable != 0 (the default): the LAST step of every hf2d_gpu_step / hf2d_gpu_step_continue call also writes them to 18
 * planes (allocated on first use) and hf2d_gpu_download_aos puts them into the records, exactly the members the reference
 * assigns on each node.  A shim that steps in chunks switches this off and on again before the last chunk of a cycle. */
int  hf2d_gpu_keep_gradients(hf2d_gpu* h, int enable);

/* Wall distance of every set, non-solid node of the uploaded grid: replaces SetMinDistanceToWall2D(J, WallNodes, x0)
 * (deeps2d_core.cpp:4783-4832; called at start-up of every Navier-Stokes run, hf2d_start.cpp:127, :171, :298 -- an
 * O(nodes x wall nodes) loop on the host).  wall_ij = the reference's WallNodes array ((i,j) int pairs in GetWallNodes
 * order, :2081-2104 -- the order decides ties: the LAST of several equidistant wall nodes wins), x0 = the sub-domain
 * offset the reference passes (0 for the whole grid).  The device l_min plane the step reads is updated in place.
 * Host outputs (each nx*ny, x-major, may be NULL):
 *   status  0 = node not visited (unset or solid: its members stay as they are), 1 = computed,
 *           2 = Tg != 0 and p == 0: the reference turns such a node into a solid one (SetCond2D(CT_SOLID_2D), :4797)
 *               -- the caller does that on its record and uploads again before stepping;
 *   l_min   FlowNode2D::l_min of status-1 nodes (other entries: the plane as uploaded, 0 without a plane);
 *   i_wall, j_wall  the nearest wall node of status-1 nodes, -1 where the reference assigns none (member unchanged). */
int  hf2d_gpu_set_min_distance_to_wall(hf2d_gpu* h, const int* wall_ij, int n_wall, double x0,
                                       double* l_min, int* i_wall, int* j_wall, unsigned char* status);
/* The same for one X-slab of a larger grid (one handle per GPU): wall_ij and the returned i_wall hold GLOBAL columns,
 * i_off = global column of the slab's column 0 (its left ghost column where it has one), nx_total = columns of the whole
 * grid.  x is formed from the global index, so every slab computes bit for bit what the undivided call computes for its
 * nodes (the reference's MPI build passes x0 = offset * dx instead, hf2d_start.cpp:171, whose rounding differs). */
int  hf2d_gpu_set_min_distance_to_wall_slab(hf2d_gpu* h, const int* wall_ij, int n_wall, int i_off, int nx_total,
                                            double* l_min, int* i_wall, int* j_wall, unsigned char* status);

/* The same pass for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  NO COUNTERPART IN THE REFERENCE.  Allocates nothing --
 * no cudaMalloc / cudaFree, no allocating library routine: the wall list is staged in a file-scope device array of
 * HF2D_NOALLOC_MAX_WALLS = 16384 (i, j) pairs (128 KB that every GPU loading the library holds until exit); longer lists are
 * refused with HF2D_EINVAL.  wall_ij is HOST memory as above; the four result planes are DEVICE buffers of the caller, nx*ny
 * entries each, none NULL, same contents as the host planes of hf2d_gpu_set_min_distance_to_wall (bit for bit); the device
 * l_min plane of the handle is updated in place.  One array per GPU shared by all calls: it keeps the last call's list
 * (each call overwrites what its kernel reads), and calls that OVERLAP on one GPU race on it -- the call returns after its
 * kernel has finished, so calls issued one after the other are safe.  The first call on a device resolves the array's address
 * (cudaGetSymbolAddress); make one call before a region in which even that is unwelcome. */
#define HF2D_NOALLOC_MAX_WALLS 16384
int  hf2d_gpu_wall_distance_noalloc(hf2d_gpu* h, const int* wall_ij, int n_wall, double x0,
                                    double* dev_l_min, int* dev_i_wall, int* dev_j_wall, unsigned char* dev_status);

/* Raw residual accumulators of the last step that reduced them -- RMS(k,ii), sumDiv(k,ii), iRMS(k,ii) and the largest
 * |dS/S| of ONE sub-domain (deeps2d_core.cpp:1122-1153), HF2D_NEQ entries each, any pointer may be NULL.  A multi-GPU
 * driver all-reduces them over its X-slabs (SUM, SUM, SUM, MAX) and finishes the norms once (:1539-1574 with a single
 * sub-domain), which is what hf2d_step_report::sum_RMS of the undivided grid holds.  Replaces the Var_pack gather of the
 * reference's MPI build (:1402-1463). */
int  hf2d_gpu_residual_sums(hf2d_gpu* h, double rms[HF2D_NEQ], double sum_div[HF2D_NEQ],
                            unsigned long long count[HF2D_NEQ], double dd_max[HF2D_NEQ]);

/* y+ of every set, non-solid node: replaces Recalc_y_plus(J, WallNodes) (deeps2d_core.cpp:2364-2386; start-up,
 * hf2d_start.cpp:299; once per output cycle on rank 0 of the MPI build, :1723-1726).  i_wall, j_wall: the nearest wall node
 * of each node (nx*ny ints each, x-major: the members SetMinDistanceToWall2D / hf2d_gpu_set_min_distance_to_wall left).
 * The wall nodes' dUdy, dVdx are those of the last step (or of the records as uploaded, before any step).  The device
 * y_plus plane the step reads is updated in place; y_plus (nx*ny doubles, may be NULL) receives the plane -- entries of
 * unset / solid nodes are the plane as uploaded.  HF2D_EINVAL if some node's (i_wall, j_wall) is not a wall node (a deck
 * without walls: the reference then reads node (0,0)'s gradients, which are not kept); y+ of those nodes is unchanged.
 * i_wall holds LOCAL columns of this handle's grid: the output of hf2d_gpu_set_min_distance_to_wall_slab (GLOBAL columns) is
 * not accepted here -- a host that cuts the grid into slabs refreshes y+ on the undivided grid before it uploads the slabs
 * (the reference does it on rank 0 only as well, deeps2d_core.cpp:1723-1726). */
int  hf2d_gpu_recalc_y_plus(hf2d_gpu* h, const int* i_wall, const int* j_wall, double* y_plus);

/* The body of the reference's Tecplot snapshot -- SaveData2D (libDEEPS2D/deeps2d_core.cpp:2589-2673, called through
 * DataSnapshot :2494-2516 at the end of every output cycle, :1773-1780): one line of 16 "%g" columns per node, j outer, i inner
 * -- formatted on the device from the current state, byte for byte what the host writer prints from the downloaded J (67 M
 * lines, ~7 GB of text at 16k x 4k).  type = the reference's ioMode (0 append / Tecplot, 1 rewrite: a blank line after each row);
 * dx_out, dy_out = (dx*MaxX)/(MaxX-1), (dy*MaxY)/(MaxY-1) as the caller's compiler evaluates them (:2618-2619).  The two
 * title lines (:2607-2616) stay the caller's.  Covers is_Cx_calc == 0 and is_p_asterisk_out == 0 on an undivided grid whose record
 * image is resident in HBM; otherwise HF2D_EINVAL / HF2D_ESTATE and the caller runs the host writer.
 *   hf2d_gpu_tecplot_format(...)  -> *bytes = length of the text, kept on the device;
 *   hf2d_gpu_tecplot_fetch(...)   -> copies it into host_text (bytes long) and releases it. */
int  hf2d_gpu_tecplot_format(hf2d_gpu* h, const hf2d_node_layout* layout, int type, double dx_out, double dy_out, size_t* bytes);
int  hf2d_gpu_tecplot_fetch(hf2d_gpu* h, char* host_text, size_t bytes);

/* ---------------------------------------------------------------------------------------------------------------------
 * X-slab multi-GPU driver (csrc/hf2d_slab.inc).  Replaces, for the GPU path, the reference's multi-rank contract:
 *   DEEPS2D_Run(o_stream, pJ, pC, rank, last_rank, x0)          libDEEPS2D/deeps2d_core.hpp:206-212, call hf2d_start.cpp:285
 *   sub-domains cut from J by active-node count                 ScanArea, deeps2d_core.cpp:2153, 2207-2216; hf2d_start.cpp:149-175
 *   ghost-column exchange, dt all-reduce, Var_pack gather       deeps2d_core.cpp:1336-1400, 806-809, 1402-1463
 * One hf2d_slab per GPU.  Ghost columns travel by direct stores into the neighbour's HBM over NVLink (peer memory), each hop
 * one pack+store+signal kernel and one wait+unpack kernel; no host round trip, no library collective on the step path.
 * Usage (one process, N GPUs):  hf2d_slab_cuts -> N x hf2d_slab_create -> hf2d_slab_connect_local -> N x hf2d_slab_upload(J)
 *                               -> hf2d_slab_run(all, N, nsteps, 1, &rep) per output cycle -> N x hf2d_slab_download(J).
 *       (one process per GPU):  hf2d_slab_create(rank) -> hf2d_slab_ipc_handle -> all-gather the handles by any side channel
 *                               -> hf2d_slab_connect_ipc -> upload -> hf2d_slab_run(&mine, 1, ...) on every rank.
 * Results equal the single-GPU run of the undivided grid bit for bit (seams are exact: DESIGN.md section 5). */
typedef struct hf2d_slab hf2d_slab;
/* cuts[0..world]: slab r owns global columns [cuts[r], cuts[r+1]); (nearly) equal ACTIVE nodes per slab, every column owned */
int  hf2d_slab_cuts(const void* J, const hf2d_node_layout* layout, int nx, int ny, int world, int min_cols, int* cuts);
/* cfg_global describes the UNDIVIDED grid (nx = MaxX of the whole grid; rank / n_ranks are ignored) */
int  hf2d_slab_create(const hf2d_cfg* cfg_global, int device, int rank, int world, const int* cuts, hf2d_slab** out);
int  hf2d_slab_destroy(hf2d_slab* s);
hf2d_gpu* hf2d_slab_solver(hf2d_slab* s);                  /* the slab's single-GPU handle (tile counts, timing, field pointers ...) */
int  hf2d_slab_columns(hf2d_slab* s, int* c0, int* c1, int* ghost_left, int* ghost_right);
size_t hf2d_slab_ipc_handle_bytes(void);                   /* 64 */
int  hf2d_slab_ipc_handle(hf2d_slab* s, void* out);        /* this rank's mailbox, exportable to other processes */
int  hf2d_slab_connect_ipc(hf2d_slab* s, const void* handles /* world x hf2d_slab_ipc_handle_bytes(), rank order */);
int  hf2d_slab_connect_local(hf2d_slab* const* all, int world);   /* every rank in this process: peer access + raw pointers */
/* J_global = the undivided host image; each slab reads its column range (+ ghost columns) and writes back ONLY its owned columns */
int  hf2d_slab_upload(hf2d_slab* s, const void* J_global, const hf2d_node_layout* layout, int pinned);
int  hf2d_slab_download(hf2d_slab* s, void* J_global, const hf2d_node_layout* layout, int pinned);
/* the same with the slab's own image (ghost columns included) for callers that never hold the undivided grid */
int  hf2d_slab_upload_local(hf2d_slab* s, const void* J_slab, const hf2d_node_layout* layout, int pinned);
int  hf2d_slab_download_local(hf2d_slab* s, void* J_slab, const hf2d_node_layout* layout, int pinned);
/* nsteps steps on the slabs this process drives (all ranks, or one); new_cycle as hf2d_gpu_step (1) / _step_continue (0).
 * The report is the undivided grid's (whole-grid residual norms; dt, time global). */
int  hf2d_slab_run(hf2d_slab* const* local, int n_local, int nsteps, int new_cycle, hf2d_step_report* report);
int  hf2d_slab_read_node(hf2d_slab* const* local, int n_local, int i_global, int j, double* p, double* Tg);
/* hf2d_gpu_tecplot_format / _fetch for a grid cut into slabs, all driven by this process: every slab formats the lines of its owned
 * columns on its own GPU, the host interleaves the row segments.  host_text == NULL: only *bytes is set; call again with a buffer. */
int  hf2d_slab_tecplot(hf2d_slab* const* local, int n_local, const hf2d_node_layout* layout, int type, double dx_out, double dy_out,
                       char* host_text, size_t* bytes);

#ifdef __cplusplus
}
#endif
#endif /* HF2D_GPU_H */
