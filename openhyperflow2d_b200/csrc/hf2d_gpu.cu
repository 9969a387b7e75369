// hf2d_gpu.cu -- C ABI of the B200-native DEEPS2D step (include/hf2d_gpu.h) and the host side of the step loop.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -shared -Xcompiler -fPIC
//        (-fmad=false: the parity contract is the x86-64 reference built without FMA contraction).
#include "../../include/hf2d_gpu.h"
#include "hf2d_dev.cuh"
#include "hf2d_stage1.cuh"
#include "hf2d_point.cuh"
#include "hf2d_stage1_plain.cuh"
#include "hf2d_stage1_tma.cuh"
#include "hf2d_flux.cuh"
#include "hf2d_heat.cuh"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <chrono>
#include <vector>

static thread_local char g_err[512] = "";
static int fail(int code, const char* fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
    return code;
}
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(HF2D_ECUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); } while (0)

struct hf2d_gpu {
    hf2d_cfg cfg;
    int device;
    cudaStream_t stream;
    cudaStream_t stream0;           // the stream created with the handle
    cudaStream_t stream_b;          // side stream: the general (boundary) tiles run beside the plain ones
    cudaEvent_t ev_fork, ev_join;
    cudaStream_t stream_c;          // third lane: the general tiles of hf2d_gpu_step_part(S2_INNER) beside the plain ones
    cudaEvent_t ev_fork_c, ev_join_c;
    bool own_stream;
    cudaEvent_t ev0, ev1;
    Params P;                       // device pointers + configuration, passed by value to the kernels
    std::vector<void*> allocs;
    std::multimap<size_t, void*> cache;     // blocks of the previous upload, by byte size (dalloc)
    std::map<void*, size_t> alloc_bytes;
    DevTable* d_tab;
    StepState* d_st;
    StepState* h_st;                // pinned mirror
    bool uploaded;
    int n_used_eq;                  // 7, 8 or 9: highest equation any node advances
    // boundary lists (indices into spec[])
    std::vector<SpecNode> h_spec;
    std::vector<int> h_spec_idx_nodes;
    int *d_nlist, *d_wall_list;
    int n_nlist, n_wall;
    std::vector<int> level_first, level_count;
    bool chain_mode;                // Neumann dependencies are linear chains: one pass + chain walk instead of level launches
    int n_chains; int *d_chain_start, *d_chain_src; ChainElem* d_chain_el;
    int n_chains_local;             // chains [0, n_chains_local) do not touch the first owned column: no hand-off needed
    int* d_nlist_seam; int n_nlist_seam;   // Neumann-list nodes of the first owned column of a slab with a left neighbour
    bool prepass_local_done;
    bool ahead;                     // between hf2d_gpu_step_part AHEAD and S2_FINISH: pre-pass / halo phase 2 address the next step's planes
    long long n_active;
    unsigned last_iter_cycle;       // last_iter of the reference (fixed during one cycle)
    uint64_t launches;
    void* d_stage; size_t stage_bytes;
    void* d_image; size_t image_bytes;   // the raw record image, resident when it fits (hf2d_io.inc)
    char* d_text; size_t text_bytes;     // formatted Tecplot body waiting to be fetched (hf2d_tecplot.inc)
    double bytes_per_cell;
    // tiles (hf2d_point.cuh): TI columns x TILE_H rows, one warp each
    int TI, n_sb, n_tx;
    int *d_tiles_plain, *d_tiles_gen; int n_plain, n_gen;
    int n_plain_edge, n_gen_edge;   // the first n_*_edge tiles of each list touch a ghost column (X-slab seams): hf2d_gpu_step_part
    bool s2_pending;                // Stage-2 parts queued, planes not yet swapped (hf2d_gpu_step_part S2_FINISH does)
    int plain_ttc;                  // TurbType word of the plain tiles (-1: no plain kernel for this grid)
    // optional per-kernel timing of the two plain-tile launches (bench.py roofline): event pairs on the main stream
    bool timing; std::vector<cudaEvent_t> tev[3][2]; int tn[3];
    bool prepass_done;              // hf2d_gpu_step_prepass already ran for the current step
    double global_time; bool have_global_time;   // GlobalTime of the running cycle (hf2d_gpu_set_global_time): FlowNode2D::time at download
    int rms_every;                  // residual norms are reduced on steps with iter % rms_every == 0 and on the last step of a call
    // TMA tile staging of K1a (hf2d_stage1_tma.cuh): tensor maps of the per-equation planes; off when the grid cannot be addressed
    K1aMaps maps; K1aMaps* d_maps; bool tma_ok; bool k1t_all_big; const double* S_half0; const double* dSy8_half0;
    bool keep_grads;                // hf2d_gpu_keep_gradients: the last step of every step call writes the 18 gradient planes
    bool grads_this_step, grads_valid;
};

extern "C" const char* hf2d_gpu_last_error(void) { return g_err; }
extern "C" int hf2d_gpu_device_count(void) { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) return 0; return n; }

// Device allocations of an upload.  A re-upload (every output cycle of a host that keeps nothing on the GPU, the e2e leg of
// bench.py, the shim's start-up passes) asks for the same sizes again: the previous upload's blocks wait in `cache` (by byte
// size) and are handed out again instead of a cudaFree + cudaMalloc pair each; what is left over is freed after the upload.
template <class T> static int dalloc(hf2d_gpu* h, T** p, size_t count, bool zero = true) {
    void* q = nullptr;
    const size_t bytes = std::max<size_t>(count, 1) * sizeof(T);
    auto it = h->cache.find(bytes);
    if (it != h->cache.end()) { q = it->second; h->cache.erase(it); }
    else CK(cudaMalloc(&q, bytes));
    h->alloc_bytes[q] = bytes;
    if (zero) CK(cudaMemsetAsync(q, 0, std::max<size_t>(count, 1) * sizeof(T), h->stream));
    h->allocs.push_back(q);
    *p = (T*)q;
    return 0;
}

static int create_impl(const hf2d_cfg* cfg, int device, hf2d_gpu* h);
extern "C" int hf2d_gpu_destroy(hf2d_gpu* h);
extern "C" int hf2d_gpu_create(const hf2d_cfg* cfg, int device, hf2d_gpu** out) {
    if (!cfg || !out) return fail(HF2D_EINVAL, "null argument");
    if (cfg->nx < 1 || cfg->ny < 1) return fail(HF2D_EINVAL, "bad grid %d x %d", cfg->nx, cfg->ny);
    if (cfg->bFF < 0 || cfg->bFF > 5) return fail(HF2D_EINVAL, "BFF %d is not implemented by the reference step (0..5)", cfg->bFF);
    for (int t = 0; t < HF2D_NTABLES; t++)
        if (cfg->tab[t].n > MAX_TAB_ROWS || cfg->tab[t].n < 0) return fail(HF2D_EINVAL, "table %d has %d rows (max %d)", t, cfg->tab[t].n, MAX_TAB_ROWS);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(HF2D_ECUDA, "no CUDA device: this library has no CPU fallback");
    if (device < 0 || device >= ndev) return fail(HF2D_EINVAL, "device %d out of range (%d devices)", device, ndev);
    CK(cudaSetDevice(device));
    hf2d_gpu* h = new hf2d_gpu();        // (value-initialised: every handle member is null until it is created)
    const int rc = create_impl(cfg, device, h);
    if (rc) { const std::string msg = g_err; hf2d_gpu_destroy(h); snprintf(g_err, sizeof(g_err), "%s", msg.c_str()); return rc; }   // no leak on an early error
    *out = h;
    return HF2D_OK;
}
static int create_impl(const hf2d_cfg* cfg, int device, hf2d_gpu* h) {
    h->cfg = *cfg; h->device = device; h->uploaded = false; h->launches = 0; h->d_stage = nullptr; h->stage_bytes = 0; h->d_image = nullptr; h->image_bytes = 0;
    h->d_nlist = h->d_wall_list = nullptr; h->n_nlist = h->n_wall = 0; h->n_active = 0;
    for (int t = 0; t < HF2D_NTABLES; t++) { h->cfg.tab[t].x = nullptr; h->cfg.tab[t].y = nullptr; }
    CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    h->stream0 = h->stream; h->own_stream = true;
    CK(cudaEventCreate(&h->ev0)); CK(cudaEventCreate(&h->ev1));
    {   // side stream: few, latency-bound launches (boundary tiles, seam work) that must not queue behind the bulk kernel's blocks
        int lo = 0, hi = 0; CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CK(cudaStreamCreateWithPriority(&h->stream_b, cudaStreamNonBlocking, hi));
        CK(cudaStreamCreateWithPriority(&h->stream_c, cudaStreamNonBlocking, hi));
        CK(cudaEventCreateWithFlags(&h->ev_fork_c, cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&h->ev_join_c, cudaEventDisableTiming));
    }
    CK(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
    std::vector<DevTable> tabs(HF2D_NTABLES);
    for (int t = 0; t < HF2D_NTABLES; t++) {
        memset(&tabs[t], 0, sizeof(DevTable));
        tabs[t].n = cfg->tab[t].n;
        for (int r = 0; r < cfg->tab[t].n; r++) { tabs[t].x[r] = cfg->tab[t].x[r]; tabs[t].y[r] = cfg->tab[t].y[r]; }
        {   // look-up class (hf2d_dev.cuh TAB_*)
            const int n = tabs[t].n; const double* x = tabs[t].x;
            bool asc = n >= 3;
            for (int r = 1; r < n && asc; r++) asc = x[r - 1] < x[r];
            tabs[t].cls = (n >= 2 && (n == 2 || x[0] >= x[n - 1])) ? TAB_ENDS : asc ? TAB_ASCENDING : TAB_SCAN;
        }
    }
    if (dalloc(h, &h->d_tab, HF2D_NTABLES)) return HF2D_ECUDA;
    CK(cudaMemcpyAsync(h->d_tab, tabs.data(), sizeof(DevTable) * HF2D_NTABLES, cudaMemcpyHostToDevice, h->stream));
    if (dalloc(h, &h->d_st, 1)) return HF2D_ECUDA;
    CK(cudaMallocHost((void**)&h->h_st, sizeof(StepState)));
    CK(cudaStreamSynchronize(h->stream));
    Params& P = h->P;
    memset(&P, 0, sizeof(P));
    P.nx = cfg->nx; P.ny = cfg->ny; P.N = (long long)cfg->nx * cfg->ny;
    // sweep range of this slab (deeps2d_core.cpp:831-840): ghost columns are not swept
    P.i_lo = (cfg->n_ranks > 1 && cfg->rank > 0) ? 1 : 0;
    P.i_hi = (cfg->n_ranks > 1 && cfg->rank < cfg->n_ranks - 1) ? cfg->nx - 1 : cfg->nx;
    P.dx = cfg->dx; P.dy = cfg->dy; P.CFL = cfg->CFL; P.beta0 = cfg->beta0; P.nrbc_beta0 = cfg->nrbc_beta0;
    { const double dx_1 = 1.0 / cfg->dx, dy_1 = 1.0 / cfg->dy; P.dxn[0] = dx_1 / 1.0; P.dxn[1] = dx_1 / 2.0; P.dyn[0] = dy_1 / 1.0; P.dyn[1] = dy_1 / 2.0; }
    P.bFF = cfg->bFF; P.ns = cfg->problem_type == 1; P.axi = cfg->flow_type == 1;
    for (int q = 0; q < 4; q++) { P.Hu[q] = cfg->Hu[q]; P.Rgas[q] = cfg->Rgas[q]; }
    P.TurbStartIter = cfg->TurbStartIter; P.tem = cfg->TurbExtModel; P.isTurbulenceReset = cfg->isTurbulenceReset;
    P.SigW = cfg->SigW; P.SigF = cfg->SigF; P.delta_bl = cfg->delta_bl; P.turb_I = cfg->turb_intensity;
    P.isAdiabaticWall = cfg->isAdiabaticWall; P.isAlternateRMS = cfg->isAlternateRMS;
    P.K0 = cfg->K0; P.dt_rule = cfg->dt_rule;
    // pow(C_mu, 3/4) is a constant of the model: evaluate it with the host libm the reference links against
    P.pow_cmu_34[0] = pow(0.09, 3. / 4.); P.pow_cmu_34[1] = pow(0.0845, 3. / 4.);
    P.tab = h->d_tab; P.st = h->d_st;
    h->last_iter_cycle = cfg->last_iter;
    h->rms_every = cfg->NOutStep; h->prepass_done = false; h->timing = false; h->tn[0] = h->tn[1] = h->tn[2] = 0;
    h->TI = 16; h->n_sb = h->n_tx = 0; h->d_tiles_plain = h->d_tiles_gen = nullptr; h->n_plain = h->n_gen = 0; h->n_plain_edge = h->n_gen_edge = 0; h->s2_pending = false; h->n_chains_local = 0; h->d_nlist_seam = nullptr; h->n_nlist_seam = 0; h->prepass_local_done = false; h->ahead = false; h->plain_ttc = -1; h->tma_ok = false; h->k1t_all_big = false; h->d_maps = nullptr; h->S_half0 = nullptr; h->dSy8_half0 = nullptr; h->keep_grads = true; h->grads_this_step = false; h->grads_valid = false;
    return HF2D_OK;
}

extern "C" int hf2d_gpu_destroy(hf2d_gpu* h) {
    if (!h) return HF2D_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (void* p : h->allocs) cudaFree(p);
    for (auto& c : h->cache) cudaFree(c.second);
    if (h->d_stage) cudaFree(h->d_stage);
    if (h->d_image) cudaFree(h->d_image);
    if (h->d_text) cudaFree(h->d_text);
    if (h->h_st) cudaFreeHost(h->h_st);
    // (a handle whose creation failed half-way holds nulls)
    for (cudaEvent_t e : {h->ev0, h->ev1, h->ev_fork_c, h->ev_join_c, h->ev_fork, h->ev_join}) if (e) cudaEventDestroy(e);
    for (int a = 0; a < 3; a++) for (int b = 0; b < 2; b++) for (cudaEvent_t e : h->tev[a][b]) cudaEventDestroy(e);
    for (cudaStream_t st : {h->stream0, h->stream_b, h->stream_c}) if (st) cudaStreamDestroy(st);
    cudaGetLastError();
    delete h;
    return HF2D_OK;
}

static bool plain_ttc_supported(int ttc);
// the 18 gradient planes exist only for callers that asked for them (hf2d_gpu_keep_gradients)
static int ensure_grad_planes(hf2d_gpu* h) {
    if (h->P.grad || !h->P.ns) return 0;
    if (dalloc(h, &h->P.grad, (size_t)G_COUNT * h->P.N)) return HF2D_ECUDA;
    return 0;
}
// ---- tensor maps for the TMA unit (driver entry point fetched through the runtime: no link against libcuda) ----
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled tma_encoder() {
    static PFN_encodeTiled fn = nullptr; static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr; cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) fn = (PFN_encodeTiled)p;
        else cudaGetLastError();
    }
    return fn;
}
// planes of nx*ny doubles, `nplanes` of them N apart (rank 3) or one plane (rank 2); box = rows x cols (x 1 plane)
static bool tma_map(CUtensorMap* m, const double* base, int nx, int ny, int nplanes, int box_rows, int box_cols) {
    PFN_encodeTiled enc = tma_encoder();
    if (!enc || !base) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)ny, (cuuint64_t)nx, (cuuint64_t)nplanes};
    const cuuint64_t strides[2] = {(cuuint64_t)ny * 8, (cuuint64_t)nx * (cuuint64_t)ny * 8};
    const cuuint32_t box[3] = {(cuuint32_t)box_rows, (cuuint32_t)box_cols, 1};
    const cuuint32_t es[3] = {1, 1, 1};
    const int rank = nplanes > 1 ? 3 : 2;
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, (cuuint32_t)rank, (void*)base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
// after the planes are allocated (hf2d_gpu_upload_aos): K1a's tensor maps; any failure leaves the LDG kernel in charge
static void build_tma_maps(hf2d_gpu* h) {
    const Params& P = h->P;
    h->tma_ok = false;
    const char* off = getenv("HF2D_NO_TMA");
    if (off && off[0] == '1') return;
    { const char* e = getenv("HF2D_K1T_ALL_BIG"); h->k1t_all_big = e && e[0] == '1'; }   // (tests: every equation through the full-stage launch)
    if (h->TI != K1T_TI || (P.ny & 1) || h->n_plain == 0) return;       // TMA strides are multiples of 16 bytes
    memset(&h->maps, 0, sizeof(h->maps));
    const double* s0 = P.S_in < P.S_out ? P.S_in : P.S_out;
    h->S_half0 = s0;
    bool ok = tma_map(&h->maps.S[0], s0, P.nx, P.ny, NEQ, K1T_S_ROWS, K1T_TI + 2) && tma_map(&h->maps.S[1], s0 + (long long)NEQ * P.N, P.nx, P.ny, NEQ, K1T_S_ROWS, K1T_TI + 2) &&
              tma_map(&h->maps.A, P.A, P.nx, P.ny, NEQ, 32, K1T_TI + 2) && tma_map(&h->maps.B, P.B, P.nx, P.ny, NEQ, K1T_B_ROWS, K1T_TI) &&
              tma_map(&h->maps.beta, P.beta, P.nx, P.ny, NEQ, 32, K1T_TI);
    if (ok && P.Src) ok = tma_map(&h->maps.Src, P.Src, P.nx, P.ny, NEQ, 32, K1T_TI);
    if (ok && P.F) ok = tma_map(&h->maps.F, P.F, P.nx, P.ny, NEQ, 32, K1T_TI);
    h->dSy8_half0 = nullptr;
    if (ok && P.dSy_in[8]) {
        const double* d0 = P.dSy_in[8] < P.dSy_out[8] ? P.dSy_in[8] : P.dSy_out[8];
        h->dSy8_half0 = d0;
        ok = tma_map(&h->maps.dSy[0], d0, P.nx, P.ny, 1, K1T_DS_ROWS, K1T_TI) && tma_map(&h->maps.dSy[1], d0 + P.N, P.nx, P.ny, 1, K1T_DS_ROWS, K1T_TI);
    }
    if (ok) {
        if (dalloc(h, &h->d_maps, 1)) { cudaGetLastError(); return; }
        if (cudaMemcpyAsync(h->d_maps, &h->maps, sizeof(K1aMaps), cudaMemcpyHostToDevice, h->stream) != cudaSuccess || cudaStreamSynchronize(h->stream) != cudaSuccess) { cudaGetLastError(); return; }
    }
    h->tma_ok = ok;
}
extern "C" int hf2d_gpu_keep_gradients(hf2d_gpu* h, int enable) {
    if (!h) return fail(HF2D_EINVAL, "null");
    h->keep_grads = enable != 0;
    return HF2D_OK;
}
#include "hf2d_io.inc"
#include "hf2d_step.inc"
#include "hf2d_walldist.inc"
#include "hf2d_noalloc.inc"
#include "hf2d_tecplot.inc"
