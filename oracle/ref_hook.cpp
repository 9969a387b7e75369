// TEST INFRASTRUCTURE ONLY (oracle/): never linked into the product library.
//
// ref_hook.cpp -- interposes on the reference's solver entry point so that the
// UNMODIFIED reference driver (hf2d_start.cpp: main -> InitSharedData ->
// InitDEEPS2D -> ... -> DEEPS2D_Run) can be used as a golden-vector generator.
//
// The reference's deeps2d_core.cpp is compiled with -DDEEPS2D_Run=DEEPS2D_Run_impl,
// hf2d_start.cpp is compiled as is, so main() lands here.  We then
//   1. write the FlowNode2D AoS layout (offsetof on the reference's own header),
//   2. write every scalar/table the DEEPS2D step reads (the hf2d_cfg contents),
//   3. write the raw AoS grid as it is handed to the solver  (state before step 1),
//   4. call the real DEEPS2D_Run, which runs Nmax steps per cycle and syncs the
//      raw AoS grid into the swap file <Project>.hf2d at the end of each cycle
//      (deeps2d_core.cpp:1818-1849) -- that file is the "after N steps" vector.
// An atexit handler prints GlobalTime/last_iter so the caller can verify how many
// cycles ran (the reference always leaves through exit(0), deeps2d_core.cpp:2022).
//
// Output directory: $OHF_DUMP_DIR (default ".").
#include "libDEEPS2D/deeps2d_core.hpp"
#include <unistd.h>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <string>

typedef FlowNode2D<FP, NUM_COMPONENTS> Node;

// globals of deeps2d_core.cpp that its header does not export
extern FP dt, beta0, CFL, nrbc_beta0, SigW, SigF, GlobalTime, CurrentTimePart, ExitMonitorValue, Ts0;
extern Table *CFL_Scenario, *beta_Scenario;
extern BlendingFactorFunction bFF;
extern int TurbStartIter, TurbExtModel, isTurbulenceReset, isAdiabaticWall, isAlternateRMS;
extern int Nstep, NOutStep, MonitorIndex;
extern unsigned int AddSrcStartIter, last_iter;

void DEEPS2D_Run_impl(ofstream* o_stream);

static std::string g_dir;

static void at_exit_report() {
    std::string fn = g_dir + "/final.txt";
    FILE* f = fopen(fn.c_str(), "w");
    if (!f) return;
    fprintf(f, "GlobalTime %.17g\nlast_iter %u\ndt %.17g\n", (double)GlobalTime, last_iter, (double)dt);
    fclose(f);
}

static void put_table(FILE* f, const char* name, Table* t) {
    if (t == NULL || t == InputData::GetZeroTable()) {
        fprintf(f, "table %s 0\n", name);
        return;
    }
    fprintf(f, "table %s %u", name, t->n);
    for (unsigned i = 0; i < t->n; i++) fprintf(f, " %.17g %.17g", (double)t->x[i], (double)t->y[i]);
    fprintf(f, "\n");
}

#define OFF(f, name, member, cnt) \
    fprintf(f, "field %s %zu %zu %d\n", name, (size_t)((char*)&(n0->member) - (char*)n0), sizeof(n0->member) / (cnt), (int)(cnt))

static void write_layout(const char* fn) {
    FILE* f = fopen(fn, "w");
    Node* n0 = (Node*)malloc(sizeof(Node));
    const int NE = Node::NumEq;
    fprintf(f, "node_bytes %zu\nnum_eq %d\nnum_components %d\n", sizeof(Node), NE, NUM_COMPONENTS);
    OFF(f, "S", S[0], 1);        fprintf(f, "count S %d\n", NE);
    OFF(f, "dSdx", dSdx[0], 1);  fprintf(f, "count dSdx %d\n", NE);
    OFF(f, "dSdy", dSdy[0], 1);  fprintf(f, "count dSdy %d\n", NE);
    OFF(f, "TurbType", TurbType, 1);
    OFF(f, "l_min", l_min, 1);
    OFF(f, "y_plus", y_plus, 1);
    OFF(f, "Re_local", Re_local, 1);
    OFF(f, "mu_t", mu_t, 1);
    OFF(f, "lam_t", lam_t, 1);
    OFF(f, "dkdx", dkdx, 1);
    OFF(f, "dkdy", dkdy, 1);
    OFF(f, "depsdx", depsdx, 1);
    OFF(f, "depsdy", depsdy, 1);
    OFF(f, "x", x, 1);
    OFF(f, "y", y, 1);
    OFF(f, "ix", ix, 1);
    OFF(f, "iy", iy, 1);
    OFF(f, "p", p, 1);
    OFF(f, "idXl", idXl, 1);
    OFF(f, "idYu", idYu, 1);
    OFF(f, "idXr", idXr, 1);
    OFF(f, "idYd", idYd, 1);
    OFF(f, "NGX", NGX, 1);
    OFF(f, "NGY", NGY, 1);
    OFF(f, "CT", CT, 1);
    OFF(f, "i_wall", i_wall, 1);
    OFF(f, "j_wall", j_wall, 1);
    OFF(f, "beta", beta[0], 1);  fprintf(f, "count beta %d\n", NE);
    OFF(f, "Q_conv", Q_conv, 1);
    OFF(f, "time", time, 1);
    OFF(f, "k", k, 1);
    OFF(f, "R", R, 1);
    OFF(f, "lam", lam, 1);
    OFF(f, "mu", mu, 1);
    OFF(f, "CP", CP, 1);
    OFF(f, "Diff", Diff, 1);
    OFF(f, "Tf", Tf, 1);
    OFF(f, "A", A[0], 1);        fprintf(f, "count A %d\n", NE);
    OFF(f, "B", B[0], 1);        fprintf(f, "count B %d\n", NE);
    OFF(f, "F", F[0], 1);        fprintf(f, "count F %d\n", NE);
    OFF(f, "RX", RX[0], 1);      fprintf(f, "count RX %d\n", NE);
    OFF(f, "RY", RY[0], 1);      fprintf(f, "count RY %d\n", NE);
    OFF(f, "Src", Src[0], 1);    fprintf(f, "count Src %d\n", NE);
    OFF(f, "SrcAdd", SrcAdd[0], 1); fprintf(f, "count SrcAdd %d\n", NE);
    OFF(f, "Tg", Tg, 1);
    OFF(f, "U", U, 1);
    OFF(f, "V", V, 1);
    OFF(f, "Y", Y[0], 1);        fprintf(f, "count Y %d\n", NUM_COMPONENTS + 1);
    OFF(f, "Uw", Uw, 1);
    OFF(f, "Vw", Vw, 1);
    OFF(f, "droYdx", droYdx[0], 1); fprintf(f, "count droYdx %d\n", NUM_COMPONENTS + 1);
    OFF(f, "droYdy", droYdy[0], 1); fprintf(f, "count droYdy %d\n", NUM_COMPONENTS + 1);
    OFF(f, "dUdx", dUdx, 1);
    OFF(f, "dUdy", dUdy, 1);
    OFF(f, "dVdx", dVdx, 1);
    OFF(f, "dVdy", dVdy, 1);
    OFF(f, "dTdx", dTdx, 1);
    OFF(f, "dTdy", dTdy, 1);
    OFF(f, "BGX", BGX, 1);
    OFF(f, "BGY", BGY, 1);
    free(n0);
    fclose(f);
}

static void write_cfg(const char* fn) {
    FILE* f = fopen(fn, "w");
    ChemicalReactionsModelData2D* c = &chemical_reactions;
    fprintf(f, "MaxX %u\nMaxY %u\n", MaxX, MaxY);
    fprintf(f, "dx %.17g\ndy %.17g\ndt %.17g\n", (double)dx, (double)dy, (double)dt);
    fprintf(f, "CFL %.17g\nbeta0 %.17g\nnrbc_beta0 %.17g\n", (double)CFL, (double)beta0, (double)nrbc_beta0);
    fprintf(f, "bFF %d\nProblemType %d\nFT %d\n", (int)bFF, (int)ProblemType, (int)Node::FT);
    fprintf(f, "Hu %.17g %.17g %.17g %.17g\n", (double)Node::Hu[0], (double)Node::Hu[1], (double)Node::Hu[2], (double)Node::Hu[3]);
    fprintf(f, "TurbStartIter %d\nTurbExtModel %d\nisTurbulenceReset %d\n", TurbStartIter, TurbExtModel, isTurbulenceReset);
    fprintf(f, "SigW %.17g\nSigF %.17g\ndelta_bl %.17g\n", (double)SigW, (double)SigF, (double)delta_bl);
    // FlowNodeTurbulence2D::I is private static; its only value is the in-class default
    // (hyper_flow_turbulence.hpp:133) because SetTurbulenceIntensity2D is never called.
    fprintf(f, "TurbIntensity %.17g\n", 0.005);
    fprintf(f, "isAdiabaticWall %d\nisAlternateRMS %d\n", isAdiabaticWall, isAlternateRMS);
    fprintf(f, "AddSrcStartIter %u\nlast_iter %u\n", AddSrcStartIter, last_iter);
    fprintf(f, "Nstep %d\nNOutStep %d\nMonitorIndex %d\n", Nstep, NOutStep, MonitorIndex);
    fprintf(f, "ExitMonitorValue %.17g\nGlobalTime %.17g\n", (double)ExitMonitorValue, (double)GlobalTime);
    fprintf(f, "K0 %.17g\nTf %.17g\n", (double)c->K0, (double)c->Tf);
    fprintf(f, "R_Fuel %.17g\nR_OX %.17g\nR_cp %.17g\nR_air %.17g\n", (double)c->R_Fuel, (double)c->R_OX, (double)c->R_cp, (double)c->R_air);
    fprintf(f, "H_Fuel %.17g\nH_OX %.17g\nH_cp %.17g\nH_air %.17g\n", (double)c->H_Fuel, (double)c->H_OX, (double)c->H_cp, (double)c->H_air);
    put_table(f, "CFL_Scenario", CFL_Scenario);
    put_table(f, "beta_Scenario", beta_Scenario);
    put_table(f, "Cp_Fuel", c->Cp_Fuel);   put_table(f, "Cp_OX", c->Cp_OX);
    put_table(f, "Cp_cp", c->Cp_cp);       put_table(f, "Cp_air", c->Cp_air);
    put_table(f, "lam_Fuel", c->lam_Fuel); put_table(f, "lam_OX", c->lam_OX);
    put_table(f, "lam_cp", c->lam_cp);     put_table(f, "lam_air", c->lam_air);
    put_table(f, "mu_Fuel", c->mu_Fuel);   put_table(f, "mu_OX", c->mu_OX);
    put_table(f, "mu_cp", c->mu_cp);       put_table(f, "mu_air", c->mu_air);
    fclose(f);
}

void DEEPS2D_Run(ofstream* o_stream) {
    const char* d = getenv("OHF_DUMP_DIR");
    g_dir = d ? d : ".";
    write_layout((g_dir + "/layout.txt").c_str());
    write_cfg((g_dir + "/cfg.txt").c_str());
    {
        FILE* f = fopen((g_dir + "/state0.hf2d").c_str(), "wb");
        fwrite(J->GetMatrixPtr(), 1, (size_t)J->GetMatrixSize(), f);
        fclose(f);
    }
    if (getenv("OHF_DUMP_ONLY")) _exit(0);       // bench.py --config: only the grid image is wanted, no CPU step, no writers
    const char* emv = getenv("OHF_EXIT_TIME");   // override the exit test: run until GlobalTime >= value
    if (emv) { MonitorIndex = 5; ExitMonitorValue = atof(emv); }
    const char* ns = getenv("OHF_NSTEP");        // override Nstep (the deck parser forces Nstep > NOutStep)
    if (ns) Nstep = atoi(ns);
    atexit(at_exit_report);
    DEEPS2D_Run_impl(o_stream);
}
