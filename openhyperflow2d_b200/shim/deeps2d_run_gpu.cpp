// deeps2d_run_gpu.cpp -- the reference-side shim: DEEPS2D_Run (libDEEPS2D/deeps2d_core.cpp:512-2023) with its inner
// time loop (:735-1642) replaced by libhf2d_gpu.so (include/hf2d_gpu.h).
//
// Build (oracle/Makefile, target `shim`): the reference's deeps2d_core.cpp is compiled with
// -DDEEPS2D_Run=DEEPS2D_Run_cpu, every other reference source as is, and this file supplies DEEPS2D_Run.  The result
// (oracle/_ref/ohf_gpu) is the reference program -- its own main(), input-deck parser, InitDEEPS2D, geometry
// rasterisers, Tecplot/RMS/monitor writers, swap-file sync and exit path -- running its time steps on the GPU.
// This file includes the reference's headers and is compiled only where the reference tree is present; nothing of
// the reference is copied here.  What stays on the host is exactly what the reference does once per output cycle or
// every NOutStep steps; the per-cell work is behind the C ABI.
#include "libDEEPS2D/deeps2d_core.hpp"
#include "hf2d_gpu.h"
#include <sys/time.h>
#include <unistd.h>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>

typedef FlowNode2D<FP, NUM_COMPONENTS> Node;

// file-scope globals of deeps2d_core.cpp that its header does not export (deeps2d_core.cpp:20-158)
extern FP dt, beta0, CFL, nrbc_beta0, SigW, SigF, GlobalTime, CurrentTimePart, ExitMonitorValue, Ts0;
extern Table *CFL_Scenario, *beta_Scenario;
extern BlendingFactorFunction bFF;
extern int TurbStartIter, TurbExtModel, isTurbulenceReset, isAdiabaticWall, isAlternateRMS;
extern int Nstep, NOutStep, NSaveStep, I, MonitorIndex, MonitorCondition, isRun, turb_mod_name_index;
extern unsigned int AddSrcStartIter, last_iter;
extern UArray<XCut>* XCutArray;
extern int is_Cx_calc, Cx_Flow_index, Cp_Flow_index, y_max, y_min, isOutHeatFluxX, isOutHeatFluxY;
extern FP x0_body, y0_body, dx_body, dy_body;
extern const char* RMS_Name[11];
extern char RMSFileName[255], MonitorsFileName[255], ErrFileName[255], TecPlotFileName[255];
extern ofstream *pRMS_OutFile, *pMonitors_OutFile, *pHeatFlux_OutFile;
extern UArray<Flow2D*>* Flow2DList;

// ---- one GPU or N: the same five verbs ----
// HF2D_GPUS=N (N > 1): the grid is cut into N X-slabs with equal numbers of active nodes (hf2d_slab_cuts, the rule of
// ScanArea, deeps2d_core.cpp:2153, 2207-2216), one slab per GPU (devices r mod the number of GPUs present), all driven from
// this one host thread -- what hf2d_start.cpp:149-175 + the per-rank DEEPS2D_Run(o_stream, pJ, pC, rank, last_rank, x0) of
// the MPI build (deeps2d_core.hpp:206-212, hf2d_start.cpp:285) do with one process per sub-domain.  Every slab reads its
// columns from J and writes its owned columns back into J, so the per-cycle writers below see one grid.
static hf2d_gpu* g_h = NULL;
static hf2d_slab* g_slab[64];
static int g_nslab = 0;
static int eng_upload(const hf2d_node_layout* lay, int pinned) {
    if (!g_nslab) return hf2d_gpu_upload_aos(g_h, J->GetMatrixPtr(), lay, pinned);
    for (int r = 0; r < g_nslab; r++) { const int rc = hf2d_slab_upload(g_slab[r], J->GetMatrixPtr(), lay, pinned); if (rc) return rc; }
    return 0;
}
static int eng_download(const hf2d_node_layout* lay, int pinned) {
    if (!g_nslab) return hf2d_gpu_download_aos(g_h, J->GetMatrixPtr(), lay, pinned);
    for (int r = 0; r < g_nslab; r++) { const int rc = hf2d_slab_download(g_slab[r], J->GetMatrixPtr(), lay, pinned); if (rc) return rc; }
    return 0;
}
static int eng_step(int first_of_cycle, int n, hf2d_step_report* rep) {
    if (!g_nslab) return first_of_cycle ? hf2d_gpu_step(g_h, n, rep) : hf2d_gpu_step_continue(g_h, n, rep);
    return hf2d_slab_run(g_slab, g_nslab, n, first_of_cycle, rep);
}
static void eng_keep_gradients(int on) {
    if (!g_nslab) { hf2d_gpu_keep_gradients(g_h, on); return; }
    for (int r = 0; r < g_nslab; r++) hf2d_gpu_keep_gradients(hf2d_slab_solver(g_slab[r]), on);
}
static void eng_read_node(int i, int j, double* p, double* T) {
    if (!g_nslab) hf2d_gpu_read_node(g_h, i, j, p, T); else hf2d_slab_read_node(g_slab, g_nslab, i, j, p, T);
}
static void eng_set_global_time(double t) {
    if (!g_nslab) { hf2d_gpu_set_global_time(g_h, t); return; }
    for (int r = 0; r < g_nslab; r++) hf2d_gpu_set_global_time(hf2d_slab_solver(g_slab[r]), t);
}
static void eng_destroy() {
    if (!g_nslab) { hf2d_gpu_destroy(g_h); return; }
    for (int r = 0; r < g_nslab; r++) hf2d_slab_destroy(g_slab[r]);
}

// ---- the Tecplot snapshot with its body formatted on the device (HF2D_GPU_TECPLOT=1) ----
// DataSnapshot (deeps2d_core.cpp:2494-2516) = CutFile if rewriting, OpenData (append mode), SaveData2D, close.  The two title
// lines are written here exactly as SaveData2D writes them (:2594-2616); the 16-column body comes from
// hf2d_gpu_tecplot_format / _fetch.  Falls back to the reference's own DataSnapshot where the formatter does not apply
// (is_Cx_calc, is_p_asterisk_out, no resident image).
extern int is_p_asterisk_out;
static int g_gpu_tecplot = 0;
static void snapshot(char* filename, WRITE_MODE mode, const hf2d_node_layout* lay) {
    size_t bytes = 0;
    const FP dx_out = (dx * MaxX) / (MaxX - 1), dy_out = (dy * MaxY) / (MaxY - 1);       // :2618-2619
    if (!g_gpu_tecplot || is_Cx_calc || is_p_asterisk_out ||
        (g_nslab ? hf2d_slab_tecplot(g_slab, g_nslab, lay, (int)mode, dx_out, dy_out, NULL, &bytes)
                 : hf2d_gpu_tecplot_format(g_h, lay, (int)mode, dx_out, dy_out, &bytes))) { DataSnapshot(filename, mode); return; }
    char* text = (char*)malloc(bytes ? bytes : 1);
    if (!text || (g_nslab ? hf2d_slab_tecplot(g_slab, g_nslab, lay, (int)mode, dx_out, dy_out, text, &bytes)
                          : hf2d_gpu_tecplot_fetch(g_h, text, bytes))) { free(text); DataSnapshot(filename, mode); return; }
    if (mode) CutFile(filename);
    ofstream* out = OpenData(filename);
    char t1[1024] = {0}, t2[256] = {0};
    snprintf(t1, 1024, "VARIABLES = X, %s, U, V, T, p, Rho, Y_fuel, Y_ox, Y_cp, Y_i, %s, Mach, l_min, y+, Cp\n",
             FlowNode2D<FP, NUM_COMPONENTS>::FT == 1 ? "R" : "Y", "mu_t/mu");
    snprintf(t2, 256, "ZONE T=\"Time: %g sec.\" I= %i J= %i F=POINT\n", GlobalTime, MaxX, MaxY);
    *out << t1 << t2;
    out->write(text, (std::streamsize)bytes);
    out->close();
    free(text);
}

static void die(ofstream* f, const char* what) {
    *f << "\nhf2d_gpu: " << what << ": " << hf2d_gpu_last_error() << "\n" << flush;
    Abort_OpenHyperFLOW2D();
}

static void set_table(hf2d_table* t, Table* src) {
    if (src == NULL || src == InputData::GetZeroTable() || src->n == 0) { t->n = 0; t->x = t->y = NULL; return; }
    t->n = (int)src->n;
    double* x = new double[src->n]; double* y = new double[src->n];
    for (unsigned i = 0; i < src->n; i++) { x[i] = src->x[i]; y[i] = src->y[i]; }
    t->x = x; t->y = y;
}

#define OFFS(m) L.m = (size_t)((char*)&(n0->m) - (char*)n0)
static hf2d_node_layout node_layout() {
    hf2d_node_layout L; memset(&L, 0, sizeof(L));
    Node* n0 = (Node*)malloc(sizeof(Node));
    L.node_bytes = sizeof(Node);
    L.S = (size_t)((char*)&(n0->S[0]) - (char*)n0); L.dSdx = (size_t)((char*)&(n0->dSdx[0]) - (char*)n0); L.dSdy = (size_t)((char*)&(n0->dSdy[0]) - (char*)n0);
    OFFS(TurbType); OFFS(l_min); OFFS(y_plus); OFFS(mu_t); OFFS(lam_t); OFFS(dkdx); OFFS(dkdy); OFFS(depsdx); OFFS(depsdy); OFFS(y); OFFS(p);
    OFFS(idXl); OFFS(idYu); OFFS(idXr); OFFS(idYd); OFFS(CT);
    L.beta = (size_t)((char*)&(n0->beta[0]) - (char*)n0);
    OFFS(Q_conv); OFFS(time); OFFS(k); OFFS(R); OFFS(lam); OFFS(mu); OFFS(CP); OFFS(Diff); OFFS(Tf);
    L.A = (size_t)((char*)&(n0->A[0]) - (char*)n0); L.B = (size_t)((char*)&(n0->B[0]) - (char*)n0); L.F = (size_t)((char*)&(n0->F[0]) - (char*)n0);
    L.Src = (size_t)((char*)&(n0->Src[0]) - (char*)n0); L.SrcAdd = (size_t)((char*)&(n0->SrcAdd[0]) - (char*)n0);
    OFFS(Tg); OFFS(U); OFFS(V);
    L.Y = (size_t)((char*)&(n0->Y[0]) - (char*)n0);
    OFFS(Uw); OFFS(Vw);
    L.droYdx = (size_t)((char*)&(n0->droYdx[0]) - (char*)n0); L.droYdy = (size_t)((char*)&(n0->droYdy[0]) - (char*)n0);
    OFFS(dUdx); OFFS(dUdy); OFFS(dVdx); OFFS(dVdy); OFFS(dTdx); OFFS(dTdy); OFFS(BGX); OFFS(BGY);
    free(n0);
    return L;
}

static void fill_cfg(hf2d_cfg* c) {
    memset(c, 0, sizeof(*c));
    ChemicalReactionsModelData2D* ch = &chemical_reactions;
    c->nx = (int)MaxX; c->ny = (int)MaxY; c->dx = dx; c->dy = dy; c->dt = dt;
    c->CFL = CFL; c->beta0 = beta0; c->nrbc_beta0 = nrbc_beta0; c->bFF = (int)bFF;
    c->problem_type = (int)ProblemType; c->flow_type = (int)Node::FT;
    for (int q = 0; q < 4; q++) c->Hu[q] = Node::Hu[q];
    c->TurbStartIter = TurbStartIter; c->TurbExtModel = TurbExtModel; c->isTurbulenceReset = isTurbulenceReset;
    c->SigW = SigW; c->SigF = SigF; c->delta_bl = delta_bl;
    c->turb_intensity = 0.005;   // FlowNodeTurbulence2D::I: private static, in-class default, never set (hyper_flow_turbulence.hpp:133)
    c->isAdiabaticWall = isAdiabaticWall; c->isAlternateRMS = isAlternateRMS;
    c->AddSrcStartIter = AddSrcStartIter; c->last_iter = last_iter;
    c->MonitorIndex = MonitorIndex; c->ExitMonitorValue = ExitMonitorValue;
    c->K0 = ch->K0; c->Rgas[0] = ch->R_Fuel; c->Rgas[1] = ch->R_OX; c->Rgas[2] = ch->R_cp; c->Rgas[3] = ch->R_air;
    Table* tabs[HF2D_NTABLES] = {ch->Cp_Fuel, ch->Cp_OX, ch->Cp_cp, ch->Cp_air, ch->lam_Fuel, ch->lam_OX, ch->lam_cp, ch->lam_air,
                                 ch->mu_Fuel, ch->mu_OX, ch->mu_cp, ch->mu_air, CFL_Scenario, beta_Scenario};
    for (int t = 0; t < HF2D_NTABLES; t++) set_table(&c->tab[t], tabs[t]);
    c->dt_rule = 0;              // the serial build's running minimum (deeps2d_core.cpp:803, 1326)
    c->rank = 0; c->n_ranks = 1;
    c->NOutStep = NOutStep;
}

void DEEPS2D_Run(ofstream* f_stream) {
    struct timeval start, stop, mark1, mark2;
    // --- files (deeps2d_core.cpp:663-672) ---
    snprintf(RMSFileName, 255, "RMS-%s", OutFileName);
    CutFile(RMSFileName);
    pRMS_OutFile = OpenData(RMSFileName);
    SaveRMSHeader(pRMS_OutFile);
    if (MonitorPointsArray && MonitorPointsArray->GetNumElements() > 0) {
        snprintf(MonitorsFileName, 255, "Monitors-%s", OutFileName);
        CutFile(MonitorsFileName);
        pMonitors_OutFile = OpenData(MonitorsFileName);
        SaveMonitorsHeader(pMonitors_OutFile, MonitorPointsArray);
    }
    // --- the grid goes to the GPU once; it comes back before every per-cycle writer runs ---
    hf2d_cfg cfg; fill_cfg(&cfg);
    const hf2d_node_layout lay = node_layout();
    hf2d_gpu* h = NULL;
    const char* dev = getenv("HF2D_DEVICE");
    const int n_gpus = getenv("HF2D_GPUS") ? atoi(getenv("HF2D_GPUS")) : 1;
    if (n_gpus > 1) {
        if (n_gpus > 64) die(f_stream, "HF2D_GPUS > 64");
        const int n_dev = hf2d_gpu_device_count();
        if (n_dev < 1) die(f_stream, "no CUDA device");
        int cuts[65];
        if (hf2d_slab_cuts(J->GetMatrixPtr(), &lay, (int)MaxX, (int)MaxY, n_gpus, 2, cuts)) die(f_stream, "slab cuts");
        for (int r = 0; r < n_gpus; r++)
            if (hf2d_slab_create(&cfg, (r + (dev ? atoi(dev) : 0)) % n_dev, r, n_gpus, cuts, &g_slab[r])) die(f_stream, "slab create");
        g_nslab = n_gpus;
        if (hf2d_slab_connect_local(g_slab, g_nslab)) die(f_stream, "slab connect");
        *f_stream << "hf2d_gpu: " << n_gpus << " X-slabs on " << (n_gpus < n_dev ? n_gpus : n_dev) << " GPU(s), cuts at columns";
        for (int r = 0; r <= n_gpus; r++) *f_stream << " " << cuts[r];
        *f_stream << "\n" << flush;
    } else {
        if (hf2d_gpu_create(&cfg, dev ? atoi(dev) : 0, &h)) die(f_stream, "create");
        g_h = h;
    }
    // J lives as long as the run: page-lock it once so every per-cycle download runs at PCIe speed (optional)
    const int pinned = hf2d_gpu_pin_host(J->GetMatrixPtr(), (size_t)J->GetMatrixSize()) == 0;
    // Optional second seam (INTEGRATION.md, HF2D_GPU_WALL_DISTANCE=1): SetMinDistanceToWall2D on the GPU.  main() has
    // already run the host version by the time DEEPS2D_Run is called (hf2d_start.cpp:298) -- a maintainer would drop that
    // call -- so its results are wiped from J first: what the step and the swap file see is what the GPU computed.
    const bool gpu_wall_distance = n_gpus <= 1 && getenv("HF2D_GPU_WALL_DISTANCE") && atoi(getenv("HF2D_GPU_WALL_DISTANCE")) &&
                                   ProblemType == SM_NS && WallNodes && WallNodes->GetNumElements() > 0;   // (no walls: the host's i_wall = j_wall = 0 stand)
    if (gpu_wall_distance) {
        for (int i = 0; i < MaxX; i++) for (int j = 0; j < MaxY; j++) {
            Node& nd = J->GetValue(i, j);
            if (nd.isCond2D(CT_NODE_IS_SET_2D) && !nd.isCond2D(CT_SOLID_2D)) { nd.l_min = -1.; nd.i_wall = -1; nd.j_wall = -1; }
        }
    }
    if (eng_upload(&lay, pinned)) die(f_stream, "upload");
    if (gpu_wall_distance) {
        const size_t N = (size_t)MaxX * MaxY;
        double* wl = new double[N]; int* wi = new int[N]; int* wj = new int[N]; unsigned char* ws = new unsigned char[N];
        // XY<int> is two ints (utl/umatrix2d.hpp:44-65): the array is the (i,j) pair list in GetWallNodes order
        if (hf2d_gpu_set_min_distance_to_wall(h, (const int*)WallNodes->GetArrayPtr(), (int)WallNodes->GetNumElements(), 0.0,
                                              wl, wi, wj, ws)) die(f_stream, "wall distance");
        long n_solid = 0;
        for (int i = 0; i < MaxX; i++) for (int j = 0; j < MaxY; j++) {
            const size_t n = (size_t)i * MaxY + j;
            Node& nd = J->GetValue(i, j);
            if (ws[n] == 1) { nd.l_min = wl[n]; if (wi[n] >= 0) { nd.i_wall = wi[n]; nd.j_wall = wj[n]; } }
            else if (ws[n] == 2) { nd.SetCond2D(CT_SOLID_2D); n_solid++; }        // deeps2d_core.cpp:4797
        }
        // (nodes the kernel re-typed as solid change the boundary lists: the image is uploaded again before y+)
        if (n_solid && hf2d_gpu_upload_aos(h, J->GetMatrixPtr(), &lay, pinned)) die(f_stream, "upload");
        *f_stream << "hf2d_gpu: wall distance of " << WallNodes->GetNumElements() << " wall nodes on the GPU\n" << flush;
        // y+ from the same start-up data on the device (Recalc_y_plus, hf2d_start.cpp:299), with i_wall / j_wall as J now
        // holds them.  main() has computed it on the host already: the two are compared, the device plane is what the
        // steps read.  (A deck without walls names node (0,0) as everybody's wall: refused, the host values stand.)
        for (int i = 0; i < MaxX; i++) for (int j = 0; j < MaxY; j++) {
            const size_t n = (size_t)i * MaxY + j;
            wi[n] = J->GetValue(i, j).i_wall; wj[n] = J->GetValue(i, j).j_wall;
        }
        if (hf2d_gpu_recalc_y_plus(h, wi, wj, wl) == 0) {
            long n_diff = 0;
            for (int i = 0; i < MaxX; i++) for (int j = 0; j < MaxY; j++) {
                Node& nd = J->GetValue(i, j);
                if (!nd.isCond2D(CT_NODE_IS_SET_2D) || nd.isCond2D(CT_SOLID_2D)) continue;
                const double y = wl[(size_t)i * MaxY + j];
                if (!(y == nd.y_plus)) n_diff++;
                nd.y_plus = y;
            }
            *f_stream << "hf2d_gpu: y+ on the GPU: " << n_diff << " nodes differ from the host's\n" << flush;
        } else {
            *f_stream << "hf2d_gpu: y+ left to the host (" << hf2d_gpu_last_error() << ")\n" << flush;
        }
        delete[] wl; delete[] wi; delete[] wj; delete[] ws;
        // One download contract for both image paths: every byte the step does not own comes back AS UPLOADED.  l_min,
        // y_plus, i_wall, j_wall were just written into J on the host, so J goes up once more -- the resident device
        // image (and the planes the steps read) then hold exactly what J holds now.
        if (hf2d_gpu_upload_aos(h, J->GetMatrixPtr(), &lay, pinned)) die(f_stream, "upload");
    }
    *f_stream << "hf2d_gpu: DEEPS2D time loop on the GPU (" << MaxX << "x" << MaxY << " nodes)\n" << flush;
    g_gpu_tecplot = getenv("HF2D_GPU_TECPLOT") && atoi(getenv("HF2D_GPU_TECPLOT"));
    if (g_gpu_tecplot) *f_stream << "hf2d_gpu: Tecplot snapshots formatted on the GPU\n" << flush;

    I = 0; isRun = 1;
    FP max_RMS = 0.; int k_max_RMS = -1;
    do {
        gettimeofday(&mark2, NULL);
        gettimeofday(&start, NULL);
        // one output cycle = Nstep steps; the reference looks at the residuals on steps with iter % NOutStep == 0
        hf2d_step_report rep; memset(&rep, 0, sizeof(rep));
        int done = 0;
        while (done < Nstep) {
            // the step with in-cycle index `done` is an output step iff done % NOutStep == 0: run up to and including it
            int n = (done % NOutStep == 0) ? 1 : (NOutStep - done % NOutStep);
            if (done + n > Nstep) n = Nstep - done;
            eng_keep_gradients(done + n >= Nstep);      // the records' gradient members are observable after the cycle's last step only
            const int rc = eng_step(done == 0, n, &rep);
            if (rc == HF2D_EUNSTABLE) {               // Tg < 0 (deeps2d_core.cpp:1246-1316): error snapshot + abort
                eng_download(&lay, pinned);
                *f_stream << "\nTg=" << J->GetValue(rep.err_i, rep.err_j).Tg << " K. p=" << J->GetValue(rep.err_i, rep.err_j).p
                          << " Pa dt=" << rep.dt << " sec.\nComputational unstability in node (" << rep.err_i << "," << rep.err_j << ") on iteration "
                          << rep.iter << "\n" << flush;
                DataSnapshot(ErrFileName);
                Abort_OpenHyperFLOW2D();
            } else if (rc) die(f_stream, "step");
            done += n;
            iter = (unsigned)done;
            dt = rep.dt;
            CurrentTimePart = rep.time_part;
            const unsigned it_last = (unsigned)(done - 1);          // in-cycle index of the step just finished
            max_RMS = rep.max_RMS; k_max_RMS = rep.k_max_RMS;
            if (it_last % (unsigned)NOutStep == 0) {                  // deeps2d_core.cpp:1576-1637
                if (MonitorPointsArray) {
                    for (int m = 0; m < (int)MonitorPointsArray->GetNumElements(); m++) {
                        MonitorPoint& mp = MonitorPointsArray->GetElement(m);
                        double pp = 0., tt = 0.;
                        eng_read_node((int)(mp.MonitorXY.GetX() / dx), (int)(mp.MonitorXY.GetY() / dy), &pp, &tt);
                        mp.p = pp; mp.T = tt;
                    }
                }
                if (isVerboseOutput) {
                    gettimeofday(&mark1, NULL);
                    const FP d_time = (FP)(mark1.tv_sec - mark2.tv_sec) + (FP)(mark1.tv_usec - mark2.tv_usec) * 1.e-6;
                    const FP VCOMP = d_time > 0. ? (FP)(NOutStep) / d_time : 0.;
                    memcpy(&mark2, &mark1, sizeof(mark1));
                    FP rms[HF2D_NEQ];
                    for (int k = 0; k < HF2D_NEQ; k++) rms[k] = rep.sum_RMS[k];
                    SaveRMS(pRMS_OutFile, last_iter + it_last, rms);
                    if (MonitorPointsArray && MonitorPointsArray->GetNumElements() > 0)
                        SaveMonitors(pMonitors_OutFile, GlobalTime + CurrentTimePart, MonitorPointsArray);
                    int kk = k_max_RMS;
                    if (kk == i2d_nu_t) kk += turb_mod_name_index;
                    if (kk != -1 && (MonitorIndex == 0 || MonitorIndex == 5))
                        *f_stream << "Step No " << it_last + last_iter << " maxRMS[" << RMS_Name[kk] << "]=" << (FP)(max_RMS * 100.)
                                  << " % step_time=" << (FP)d_time << " sec (" << (FP)VCOMP << " step/sec) dt=" << dt << "\n" << flush;
                    else if (MonitorIndex > 0 && MonitorIndex < 5)
                        *f_stream << "Step No " << it_last + last_iter << " maxRMS[" << RMS_Name[MonitorIndex - 1] << "]=" << (FP)(max_RMS * 100.)
                                  << " % step_time=" << (FP)d_time << " sec (" << (FP)VCOMP << " step/sec) dt=" << dt << "\n" << flush;
                    else
                        *f_stream << "Step No " << it_last + last_iter << " maxRMS[" << kk << "]=" << (FP)(max_RMS * 100.)
                                  << " % step_time=" << (FP)d_time << " sec (" << (FP)VCOMP << " step/sec) dt=" << dt << "\n" << flush;
                }
            }
        }
        // --- cycle end (deeps2d_core.cpp:1762-1849): grid back to the host image, then the reference's own writers ---
        eng_set_global_time(GlobalTime);                    // FlowNode2D::time of the swept nodes (:867): still the cycle's start value
        if (eng_download(&lay, pinned)) die(f_stream, "download");
        if (XCutArray->GetNumElements() > 0) {
            for (int i_xcut = 0; i_xcut < (int)XCutArray->GetNumElements(); i_xcut++) {
                XCut* c = XCutArray->GetElementPtr(i_xcut);
                *f_stream << "Cut(" << i_xcut + 1 << ") X=" << c->x0 << " Y=" << c->y0 << " dY=" << c->dy << " MassFlow="
                          << CalcMassFlowRateX2D(J, c->x0, c->y0, c->dy) << "  (kg/sec*m)" << endl;
            }
        }
        gettimeofday(&stop, NULL);
        *f_stream << "\nSave current results in file " << OutFileName << "...\n" << flush;
        snapshot(OutFileName, WM_REWRITE, &lay);
        if ((I / NSaveStep) * NSaveStep == I) {
            *f_stream << "Add current results to transient solution file " << TecPlotFileName << "...\n" << flush;
            snapshot(TecPlotFileName, WM_APPEND, &lay);
        }
        I++;
        const FP c_time = (FP)(stop.tv_sec - start.tv_sec) + (FP)(stop.tv_usec - start.tv_usec) * 1.e-6;
        *f_stream << "HyperFLOW/DEEPS computation cycle time=" << (FP)c_time << " sec ( average  speed " << (FP)(Nstep / c_time) << " step/sec).       \n" << flush;
        last_iter += iter;
        iter = 0;
        GlobalTime += CurrentTimePart;
        CurrentTimePart = 0.;
        if (isOutHeatFluxX) {
            char fn[255]; snprintf(fn, 255, "HeatFlux-X-%s", OutFileName); CutFile(fn);
            pHeatFlux_OutFile = OpenData(fn);
            SaveXHeatFlux2D(pHeatFlux_OutFile, J, Flow2DList->GetElement(Cp_Flow_index - 1), Ts0, y_max, y_min);
            pHeatFlux_OutFile->close();
        }
        if (isOutHeatFluxY) {
            char fn[255]; snprintf(fn, 255, "HeatFlux-Y-%s", OutFileName); CutFile(fn);
            pHeatFlux_OutFile = OpenData(fn);
            SaveYHeatFlux2D(pHeatFlux_OutFile, J, Ts0);
            pHeatFlux_OutFile->close();
        }
        if (is_Cx_calc)
            *f_stream << "\nCx = " << Calc_Cx_2D(J, x0_body, y0_body, dx_body, dy_body, Flow2DList->GetElement(Cx_Flow_index - 1))
                      << " Cy = " << Calc_Cy_2D(J, x0_body, y0_body, dx_body, dy_body, Flow2DList->GetElement(Cx_Flow_index - 1))
                      << " Fx = " << CalcXForce2D(J, x0_body, y0_body, dx_body, dy_body) << " Fy = " << CalcYForce2D(J, x0_body, y0_body, dx_body, dy_body) << endl;
        if (GasSwapData) {                                   // swap-file sync (_NO_MMAP_ build): rewrite the raw AoS image
            if (isVerboseOutput) *f_stream << "\nSync swap file for gas..." << flush;
            const char* ptr = (const char*)J->GetMatrixPtr();
            const size_t total = (size_t)J->GetMatrixSize();
            size_t len = 0;
            while (len < total) {
                const size_t one = (total - len > ((size_t)1 << 30)) ? ((size_t)1 << 30) : total - len;
                const ssize_t w = pwrite(fd_g, ptr + len, one, (off_t)len);
                if (w <= 0) break;
                len += (size_t)w;
            }
            if (len != total) *f_stream << "Error: len(" << len << ") != FileSize(" << total << ") " << endl << flush;
            *f_stream << "OK" << endl;
        }
        if (MonitorIndex < 5) MonitorCondition = (max_RMS > ExitMonitorValue) ? 1 : 0;
        else MonitorCondition = (GlobalTime < ExitMonitorValue) ? 1 : 0;
    } while (MonitorCondition);
    snapshot(OutFileName, WM_REWRITE, &lay);
    if (pinned) hf2d_gpu_unpin_host(J->GetMatrixPtr());
    eng_destroy();
    *f_stream << "\nReady. Computation finished.\n" << flush;
    Exit_OpenHyperFLOW2D();
}
