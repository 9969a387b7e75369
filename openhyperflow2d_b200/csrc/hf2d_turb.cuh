// hf2d_turb.cuh -- the gradient-dependent part of FlowNode2D::TurbModRANS2D (hyper_flow_node.hpp:602-957):
// eddy viscosity, turbulent fluxes and sources.  The S-rewriting part lives in local_update() (hf2d_local.cuh).
#pragma once
#include "hf2d_local.cuh"

struct Grads {
    double droYdx[4], droYdy[4];
    double dUdx, dUdy, dVdx, dVdy, dTdx, dTdy, dkdx, dkdy, depsdx, depsdy;
};

// turbulence outputs that live in per-equation planes 7/8; `w*` flags say which were assigned this step
struct TurbOut {
    double A7, A8, B7, B8, F7, F8, Src7, Src8;
    bool wAB7, wAB8, wF7, wF8, wSrc7, wSrc8, zSrcAdd78;
    double mu_t, lam_t;
    bool w_lam_t;
};

// the reference's pow(double,int) is __builtin_powi, expanded by g++ as x^3*x^3 with x^3=(x*x)*x (oracle/deeps2d_oracle.c)
__device__ __forceinline__ double powi6(double x) { const double x3 = (x * x) * x; return x3 * x3; }

template <class KC, class D, class LD>
__device__ __forceinline__ void turb_own(D& dv, const LD& ld, const Params& P, const StepState& st, long long m, uint32_t ct, uint32_t tt,
                                         const Local& L, const Grads& g, double yj, TurbOut& T) {
    const double rho = L.S[iRho];
    const int is_init = st.is_init, is_mu_t = st.is_mu_t, tem = P.tem;
    const double l_min = ld.lmin();
    double l = OMAX(l_min, OMIN(P.dy, P.dx)) * 0.41;
    if (is_set(tt, TCT_Prandtl_Model)) {
        const double n_0 = l_min * 0.41, A_p = 26.;
        if (tem == TEM_Prandtl) l = n_0;
        else if (tem == TEM_vanDriest) l = n_0 * (1. - exp(-P.y_plus[m] / A_p));
        else if (tem == TEM_Escudier) { if (P.delta_bl > 0.) l = OMIN(n_0, 0.09 * P.delta_bl); else l = n_0; }
        else if (tem == TEM_Klebanoff) { if (P.delta_bl > 0.) l = n_0 / sqrt(1 + 5.5 * powi6(l_min / P.delta_bl)); else l = n_0; }
        T.mu_t = rho * l * l * OMAX(fabs(g.dUdy), fabs(g.dVdx));
        T.lam_t = T.mu_t * L.CP; T.w_lam_t = true;
    } else if (is_set(tt, TCT_k_eps_Model)) {
        double C1eps = 1.44, C2eps = 1.92, C_mu = 0.09, sig_k = 1.0, sig_eps = 1.3;
        const double f1 = 1.; double f2 = 1., f_mu = 1.;
        double Rt = 0, nu, G, L_k = 0, L_eps = 0, Mt = 0;
        const double Tmp1 = g.dUdy + g.dVdx;
        const double Tmp2 = rho * l;
        double Tmp3 = g.dUdx * g.dUdx + g.dVdy * g.dVdy;
        if (KC::axi) Tmp3 += dv.div(L.Ut, yj);
        if (T.mu_t == 0) T.mu_t = rho * l * l * OMAX(fabs(g.dUdy), fabs(g.dVdx));
        G = T.mu_t * (Tmp1 * Tmp1 + 2 * Tmp3);
        // Rt and the damping functions see S[k], S[eps] as committed, before this call's boundary rewrites
        if (tem == TEM_k_eps_Chien || tem == TEM_k_eps_JL || tem == TEM_k_eps_LSY) {   // (the only variants that read Rt)
            if (L.s8c != 0.0 && L.mu != 0.0) Rt = dv.div(dv.div(L.s7c * L.s7c, L.s8c), L.mu); else Rt = 0;
        }
        if (tem == TEM_k_eps_Std) {
        } else if (tem == TEM_k_eps_Chien) {
            const double yp = P.y_plus[m];
            C1eps = 1.35; C2eps = 1.8;
            f2 = 1.0 - 0.4 / 1.8 * exp(-(Rt * Rt) / 36.0);
            f_mu = 1.0 - exp(-0.0115 * yp);
            L_k = dv.div(-2.0 * L.mu * L.s7c, Tmp2 * Tmp2);
            L_eps = dv.div(-2.0 * L.mu * L.s8c, Tmp2 * Tmp2) * exp(-yp / 2.0);
            Mt = dv.div(dv.div(1.5 * L.s7c, L.kg), L.po);
        } else if (tem == TEM_k_eps_JL) {
            f_mu = exp(-2.5 / (1.0 + Rt / 50));
        } else if (tem == TEM_k_eps_LSY) {
            f_mu = exp(-3.4 / (1.0 + Rt / 50.0) / (1.0 + Rt / 50.0));
        } else if (tem == TEM_k_eps_RNG) {
            const double nu_0 = 4.38;
            if (L.s8c != 0.) nu = dv.div(dv.sqrt(G) * L.s7c, L.s8c); else nu = 0;
            C_mu = 0.0845; C1eps = 1.42;
            C2eps = 1.68 + dv.div(C_mu * nu * nu * nu * (1.0 - dv.div(nu, nu_0)), 1. + 0.012 * nu * nu * nu);
            sig_k = 0.7194; sig_eps = 0.7194; f_mu = 1.;
        }
        const double S7 = L.S[iK], S8 = L.S[iEps];   // after the rewrites
        if (is_init) { if (S8 != 0) T.mu_t = fabs(dv.div(C_mu * f_mu * S7 * S7, S8)); }
        if (is_mu_t && S8 != 0) { const double nu_t = fabs(dv.div(C_mu * f_mu * S7 * S7, S8)); T.mu_t = OMIN(nu_t, T.mu_t); }
        if (!is_init) {
            T.A7 = S7 * L.Ut; T.A8 = S8 * L.Ut; T.B7 = S7 * L.Vt; T.B8 = S8 * L.Vt;
            const double mk = dv.div(T.mu_t, sig_k), me = dv.div(T.mu_t, sig_eps);
            const double RX7 = (L.mu + mk) * g.dkdx, RX8 = (L.mu + me) * g.depsdx;
            const double RY7 = (L.mu + mk) * g.dkdy, RY8 = (L.mu + me) * g.depsdy;
            T.A7 = T.A7 - RX7; T.A8 = T.A8 - RX8; T.B7 = T.B7 - RY7; T.B8 = T.B8 - RY8;
            T.wAB7 = T.wAB8 = true;
            if (S7 != 0.) {
                T.zSrcAdd78 = true;
                if (!is_set(tt, TCT_k_CONST)) { T.Src7 = (G - S8 * (1 + Mt) + L_k * rho); T.wSrc7 = true; }
                if (!is_set(tt, TCT_eps_CONST) && S7 != 0) {
                    T.Src8 = (dv.div(C1eps * f1 * S8, S7) * G - dv.div(C2eps * f2 * S8 * S8, S7) + L_eps * rho); T.wSrc8 = true;
                }
            }
            // TurbulenceAxisymmAddOn (:242-252)
            if (KC::axi) {
                const double FT = (double)KC::axi;
                T.F7 = FT * (L.mu + T.mu_t) * g.dkdy; T.F8 = FT * (L.mu + dv.div(T.mu_t, 1.3)) * g.depsdy; T.wF7 = T.wF8 = true;
            }
        }
    } else if (is_set(tt, TCT_SA_Model)) {
        double fv1 = 1.0;
        const double S7 = L.S[iNut];
        const bool rewritten = is_init || is_set(ct, CT_WALL_NO_SLIP) || is_set(ct, CT_WALL_LAW) || is_set(tt, TCT_k_CONST) || is_set(ct, NT_FC);
        if (!rewritten) {
            const double Cb1 = 0.1355, Cb2 = 0.622, sig = 2.0 / 3.0, _k = 0.41;
            const double Cw1 = Cb1 / (_k * _k) + (1 + Cb2) / sig;
            const double Cw2 = 0.3, Cw3 = 2.0, Cv1 = 7.1, Ct2 = 2.0, Ct4 = 0.5, C5 = 3.5;
            const double a_sound2 = L.kg * L.R * L.Tgo;     // new k, lagged R, last step's Tg (FillNode2D order)
            const double nu = dv.div(L.mu, rho);
            const double ksi = dv.div(S7, nu);
            fv1 = dv.div(ksi * ksi * ksi, ksi * ksi * ksi + Cv1 * Cv1 * Cv1);
            const double nu_hat = dv.div(dv.div(T.mu_t, rho), fv1);
            const double fv2 = 1.0 - dv.div(ksi, 1.0 + ksi * fv1);
            const double Wxy = 0.5 * (g.dVdx - g.dUdy);
            const double Omega = dv.sqrt(2.0 * Wxy * Wxy);
            double S_hat = Omega + dv.div(S7, _k * _k * l_min * l_min) * fv2;
            if (S_hat < 0.3 * Omega) S_hat = 0.3 * Omega;
            const double r = OMIN((dv.div(S7, S_hat * _k * _k * l_min * l_min)), 10.0);
            const double gg = r + Cw2 * (pow(r, 6.0) - r);
            const double fw = gg * pow((1.0 + pow(Cw3, 6.0)) / (pow(gg, 6.0) + pow(Cw3, 6.0)), 1.0 / 6.0);
            const double ft2 = Ct2 * exp(-Ct4 * ksi * ksi);
            T.A7 = S7 * L.Ut; T.B7 = S7 * L.Vt;
            const double Div_nu = (g.dkdx + g.dkdy);
            const double RX7 = dv.div((nu + S7) * g.dkdx, sig), RY7 = dv.div((nu + S7) * g.dkdy, sig);
            T.A7 = T.A7 - RX7; T.B7 = T.B7 - RY7; T.wAB7 = true;
            T.Src7 = Cb1 * (1.0 - ft2) * S_hat * S7 - (Cw1 * fw - Cb1 / (_k * _k) * ft2) * dv.div(S7, l_min) * dv.div(S7, l_min) +
                     dv.div(Cb2 * Div_nu * Div_nu, sig) - dv.div(C5 * nu_hat * nu_hat, a_sound2) * g.dUdy * g.dVdx;
            T.wSrc7 = true;
        }
        // TurbulenceAxisymmAddOn
        if (!is_init) { if (KC::axi) { T.F7 = (double)KC::axi * (L.mu / rho + S7) * g.dkdy; T.wF7 = true; } }
        else { T.F7 = 0.; T.F8 = 0.; T.wF7 = T.wF8 = true; T.Src7 = 0.; T.Src8 = 0.; T.wSrc7 = T.wSrc8 = true; }
        if (is_mu_t) { T.mu_t = OMAX(0, (rho * S7 * fv1)); T.lam_t = T.mu_t * L.CP; T.w_lam_t = true; }
    } else if (is_set(tt, TCT_SST_Model)) {
        // empty in the reference (hyper_flow_node.hpp:919-920): k-omega SST has no reference arithmetic
    } else if (is_set(tt, TCT_Integral_Model) && L.mu != 0.0) {
        // only Re_local, which nothing on the step path reads
    } else if (is_set(tt, TCT_Smagorinsky_Model)) {
        const double Cs = 0.1, _delta = sqrt(P.dx * P.dy);
        const double Wxy = 0.5 * (g.dVdx - g.dUdy);
        const double Omega = dv.sqrt(2.0 * Wxy * Wxy);
        if (is_mu_t) { T.mu_t = OMAX(0, (rho * (Cs * _delta) * (Cs * _delta) * Omega)); T.lam_t = T.mu_t * L.CP; T.w_lam_t = true; }
    }
}
