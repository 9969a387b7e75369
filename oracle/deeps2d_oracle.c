/* TEST INFRASTRUCTURE ONLY -- see deeps2d_oracle.h.  Sequential, in-place, i-outer/j-inner
 * sweep exactly as the reference's serial build performs it: no attempt at
 * speed or parallelism is made here, the point is to be obviously the same algorithm.
 * Citations are relative to /root/reference/OpenHyperFLOW2D/.                              */
#include "deeps2d_oracle.h"
#include <math.h>
#include <string.h>

/* the reference's macros, hyper_flow_node.hpp:17-23 (note: not NaN-symmetric) */
#define OMAX(a, b) (((a) > (b)) ? (a) : (b))
#define OMIN(a, b) (((a) < (b)) ? (a) : (b))

/* CondType2D bits, hyper_flow_node.hpp:63-99 */
#define CT_Rho_CONST     0x01ull
#define CT_U_CONST       0x02ull
#define CT_V_CONST       0x04ull
#define CT_T_CONST       0x08ull
#define CT_Y_CONST       0x010ull
#define CT_dRhodx_NULL   0x020ull
#define CT_dYdx_NULL     0x0200ull
#define CT_dRhody_NULL   0x0400ull
#define CT_dYdy_NULL     0x04000ull
#define CT_d2Rhodx2_NULL 0x08000ull
#define CT_d2Ydx2_NULL   0x080000ull
#define CT_d2Rhody2_NULL 0x0100000ull
#define CT_d2Ydy2_NULL   0x01000000ull
#define CT_NONREFLECTED  0x02000000ull
#define CT_WALL_NO_SLIP  0x04000000ull
#define CT_WALL_LAW      0x08000000ull
#define CT_SOLID         0x040000000ull
#define CT_NODE_IS_SET   0x080000000ull
#define NT_FC (CT_Rho_CONST | CT_U_CONST | CT_V_CONST | CT_Y_CONST | CT_T_CONST | CT_NODE_IS_SET) /* :105 */
/* TurbulenceCondType2D bits, hyper_flow_turbulence.hpp:22-61 */
#define TCT_k_CONST        0x01ull
#define TCT_eps_CONST      0x02ull
#define TCT_dkdx_NULL      0x04ull
#define TCT_dkdy_NULL      0x010ull
#define TCT_d2kdx2_NULL    0x040ull
#define TCT_d2kdy2_NULL    0x0100ull
#define TCT_k_eps_Model    0x0400ull
#define TCT_Prandtl_Model  0x0800ull
#define TCT_Integral_Model 0x01000ull
#define TCT_eps_Cmk2kXn_WALL 0x08000ull
#define TCT_SA_Model       0x010000ull
#define TCT_SST_Model      0x040000ull
#define TCT_Smagorinsky_Model 0x0200000ull
/* TurbulenceExtendedModel, hyper_flow_turbulence.hpp:63-80 */
enum { TEM_Prandtl, TEM_vanDriest, TEM_Escudier, TEM_Klebanoff, TEM_k_eps_Std, TEM_k_eps_Chien,
       TEM_k_eps_JL, TEM_k_eps_LSY, TEM_k_eps_RNG };
/* equation indices, hyper_flow_node.hpp:53-60, hyper_flow_turbulence.hpp:17-20 */
enum { iRho = 0, iRhoU = 1, iRhoV = 2, iRhoE = 3, iYfu = 4, iYox = 5, iYcp = 6, iK = 7, iNut = 7, iEps = 8 };

static int is_ct(const ohf_node* n, uint64_t m) { return (n->CT & m) == m; }        /* :279-281 */
static int is_tt(const ohf_node* n, uint64_t m) { return (n->TurbType & m) == m; }  /* turbulence.hpp:136 */

int ohf_node_bytes(void) { return (int)sizeof(ohf_node); }

/* Table::GetVal, obj_data/obj_data.cpp:1822-1860 (n==0 stands for the "zero table") */
double ohf_table_getval(const ohf_table* t, double x) {
    int i, n = t->n;
    if (n == 0) return 0.;
    if (n == 1) return t->y[0];
    if (x <= t->x[0]) {
        i = 1;
    } else if (x >= t->x[n - 1]) {
        i = n - 1;
    } else {
        for (i = 1; i < n; i++)
            if ((x >= t->x[i - 1]) && (x < t->x[i])) break;
    }
    return t->y[i] + (t->y[i - 1] - t->y[i]) * (x - t->x[i]) / (t->x[i - 1] - t->x[i]);
}

/* pow(double,int) in the reference's C++98 dialect is std::pow(double,int) = __builtin_powi, which
 * g++ -O2 expands for a constant exponent along its addition-chain table: x^6 = x^3 * x^3, x^3 = (x*x)*x
 * (gcc tree-ssa-math-opts powi_as_mults; verified bit-exact against the reference binary).  Only the Klebanoff
 * mixing length uses it (hyper_flow_node.hpp:632); the SA model calls pow(double,double). */
static double powi6(double x) {
    double x3 = (x * x) * x;
    return x3 * x3;
}

/* SetTurbulenceModel, deeps2d_core.cpp:4683-4695 */
static int add_eq(const ohf_node* n) {
    if (is_tt(n, TCT_Prandtl_Model)) return 2;
    if (is_tt(n, TCT_k_eps_Model)) return 0;
    if (is_tt(n, TCT_SA_Model)) return 1;
    return 2;
}

/* CalcChemicalReactions (CRM_ZELDOVICH), deeps2d_core.cpp:4697-4780 */
static void chemistry(ohf_node* c, const ohf_cfg* g) {
    double Y0, Yfu, Yox, Ycp, Yair;
    const ohf_table* T = g->tab;
    Yfu = c->S[iYfu] / c->S[0];
    Yox = c->S[iYox] / c->S[0];
    Ycp = c->S[iYcp] / c->S[0];
    Yair = 1. - (Yfu + Yox + Ycp);
    if (!is_ct(c, CT_Y_CONST)) {
        Y0 = 1. / (Yfu + Yox + Ycp + Yair);
        Yfu = Yfu * Y0;
        Yox = Yox * Y0;
        Ycp = Ycp * Y0;
        if (c->Tg > c->Tf) {
            if (Yox > Yfu * g->K0) {
                Yox = Yox - Yfu * g->K0;
                Yfu = 0.;
                Ycp = 1. - Yox - Yair;
            } else {
                Yfu = Yfu - Yox / g->K0;
                Yox = 0.;
                Ycp = 1. - Yfu - Yair;
            }
        }
    }
    c->R = g->Rgas[0] * Yfu + g->Rgas[1] * Yox + g->Rgas[2] * Ycp + g->Rgas[3] * Yair;
    c->CP = ohf_table_getval(&T[OHF_T_CP + 0], c->Tg) * Yfu + ohf_table_getval(&T[OHF_T_CP + 1], c->Tg) * Yox +
            ohf_table_getval(&T[OHF_T_CP + 2], c->Tg) * Ycp + ohf_table_getval(&T[OHF_T_CP + 3], c->Tg) * Yair;
    if (g->problem_type == 1) {
        c->lam = ohf_table_getval(&T[OHF_T_LAM + 0], c->Tg) * Yfu + ohf_table_getval(&T[OHF_T_LAM + 1], c->Tg) * Yox +
                 ohf_table_getval(&T[OHF_T_LAM + 2], c->Tg) * Ycp + ohf_table_getval(&T[OHF_T_LAM + 3], c->Tg) * Yair;
        c->mu = ohf_table_getval(&T[OHF_T_MU + 0], c->Tg) * Yfu + ohf_table_getval(&T[OHF_T_MU + 1], c->Tg) * Yox +
                ohf_table_getval(&T[OHF_T_MU + 2], c->Tg) * Ycp + ohf_table_getval(&T[OHF_T_MU + 3], c->Tg) * Yair;
    }
    if (Yair < 1.e-5) Yair = 0.;
    if (Ycp < 1.e-8) Ycp = 0.;
    if (Yox < 1.e-8) Yox = 0.;
    if (Yfu < 1.e-8) Yfu = 0.;
    Y0 = 1. / (Yfu + Yox + Ycp + Yair);
    Yfu = Yfu * Y0;
    Yox = Yox * Y0;
    Ycp = Ycp * Y0;
    Yair = Yair * Y0;
    c->Y[0] = Yfu;
    c->Y[1] = Yox;
    c->Y[2] = Ycp;
    c->Y[3] = Yair;
    if (!is_ct(c, CT_Y_CONST)) {
        c->S[iYfu] = fabs(Yfu * c->S[0]);
        c->S[iYox] = fabs(Yox * c->S[0]);
        c->S[iYcp] = fabs(Ycp * c->S[0]);
    }
}

/* TurbulenceAxisymmAddOn, hyper_flow_node.hpp:242-252 */
static void turb_axisymm_addon(ohf_node* c, const ohf_cfg* g, int is_init) {
    double FT = (double)g->flow_type;
    if (is_tt(c, TCT_k_eps_Model) && !is_init) {
        c->F[iK] = FT * (c->mu + c->mu_t) * c->dkdy;
        c->F[iEps] = FT * (c->mu + c->mu_t / 1.3) * c->depsdy;
    } else if (is_tt(c, TCT_SA_Model) && !is_init) {
        c->F[iNut] = FT * (c->mu / c->S[iRho] + c->S[iNut]) * c->dkdy;
    } else if (is_init) {
        c->F[iNut] = c->F[iEps] = 0.;
        c->Src[iNut] = c->Src[iEps] = 0.;
    }
}

/* FlowNode2D::TurbModRANS2D, hyper_flow_node.hpp:602-957 */
static void turb_mod_rans(ohf_node* c, const ohf_cfg* g, int is_mu_t, int is_init, int tem, double delta) {
    const double dx = g->dx, dy = g->dy;
    double l = OMAX(c->l_min, OMIN(dy, dx)) * 0.41;
    if (is_tt(c, TCT_Prandtl_Model)) {
        double n_0, A_p = 26.;
        n_0 = c->l_min * 0.41;
        if (tem == TEM_Prandtl) {
            l = n_0;
        } else if (tem == TEM_vanDriest) {
            l = n_0 * (1. - exp(-c->y_plus / A_p));
        } else if (tem == TEM_Escudier) {
            if (delta > 0.) l = OMIN(n_0, 0.09 * delta);
            else l = n_0;
        } else if (tem == TEM_Klebanoff) {
            if (delta > 0.) l = n_0 / sqrt(1 + 5.5 * powi6(c->l_min / delta));
            else l = n_0;
        }
        c->mu_t = c->S[iRho] * l * l * OMAX(fabs(c->dUdy), fabs(c->dVdx));
        c->lam_t = c->mu_t * c->CP;
    } else if (is_tt(c, TCT_k_eps_Model)) {
        double C1eps = 1.44, C2eps = 1.92, C_mu = 0.09, sig_k = 1.0, sig_eps = 1.3;
        double f1 = 1., f2 = 1., f_mu = 1.;
        double Rt, nu, G, L_k, L_eps, Mt = 0;
        double Tmp1, Tmp2, Tmp3;
        L_k = L_eps = 0;
        Tmp1 = c->dUdy + c->dVdx;
        Tmp2 = c->S[iRho] * l;
        Tmp3 = c->dUdx * c->dUdx + c->dVdy * c->dVdy;
        if (g->flow_type) Tmp3 += c->U / c->y;
        if (c->mu_t == 0) c->mu_t = c->S[iRho] * l * l * OMAX(fabs(c->dUdy), fabs(c->dVdx));
        G = c->mu_t * (Tmp1 * Tmp1 + 2 * Tmp3);
        if (c->S[iEps] != 0.0 && c->mu != 0.0) Rt = c->S[iK] * c->S[iK] / c->S[iEps] / c->mu;
        else Rt = 0;
        if (tem == TEM_k_eps_Std) {
            f_mu = 1.0; C1eps = 1.44; C2eps = 1.92; C_mu = 0.09; sig_k = 1.0; sig_eps = 1.3;
        } else if (tem == TEM_k_eps_Chien) {
            C1eps = 1.35; C2eps = 1.8; C_mu = 0.09; sig_k = 1.0; sig_eps = 1.3;
            f2 = 1.0 - 0.4 / 1.8 * exp(-(Rt * Rt) / 36.0);
            f_mu = 1.0 - exp(-0.0115 * c->y_plus);
            L_k = -2.0 * c->mu * c->S[iK] / (Tmp2 * Tmp2);
            L_eps = -2.0 * c->mu * c->S[iEps] / (Tmp2 * Tmp2) * exp(-c->y_plus / 2.0);
            Mt = 1.5 * c->S[iK] / c->k / c->p;
        } else if (tem == TEM_k_eps_JL) {
            C1eps = 1.44; C2eps = 1.92; C_mu = 0.09; sig_k = 1.0; sig_eps = 1.3;
            f_mu = exp(-2.5 / (1.0 + Rt / 50));
        } else if (tem == TEM_k_eps_LSY) {
            C1eps = 1.44; C2eps = 1.92; C_mu = 0.09; sig_k = 1.0; sig_eps = 1.3;
            f_mu = exp(-3.4 / (1.0 + Rt / 50.0) / (1.0 + Rt / 50.0));
        } else if (tem == TEM_k_eps_RNG) {
            double nu_0 = 4.38;
            if (c->S[iEps] != 0.) nu = sqrt(G) * c->S[iK] / c->S[iEps];
            else nu = 0;
            C_mu = 0.0845;
            C1eps = 1.42;
            C2eps = 1.68 + C_mu * nu * nu * nu * (1.0 - nu / nu_0) / (1. + 0.012 * nu * nu * nu);
            sig_k = 0.7194;
            sig_eps = 0.7194;
            f_mu = 1.;
        }
        if (is_init) {
            double TmpI = g->turb_intensity * sqrt(c->U * c->U + c->V * c->V + 1.e-30);
            c->S[iK] = 1.5 * TmpI * TmpI * c->S[iRho];
            c->S[iEps] = pow(C_mu, 3. / 4.) * pow(c->S[iK] / c->S[iRho], 3. / 2.) / l;
            if (c->S[iEps] != 0) c->mu_t = fabs(C_mu * f_mu * c->S[iK] * c->S[iK] / c->S[iEps]);
        }
        if (is_tt(c, TCT_k_CONST)) {
            double TmpI = g->turb_intensity * sqrt(c->U * c->U + c->V * c->V + 1.e-30);
            c->S[iK] = 1.5 * TmpI * TmpI * c->S[iRho];
        }
        if (is_tt(c, TCT_eps_CONST))
            c->S[iEps] = pow(C_mu, 3. / 4.) * pow(c->S[iK] / c->S[iRho], 3. / 2.) / l;
        if (is_tt(c, TCT_eps_Cmk2kXn_WALL))
            c->S[iEps] = pow(C_mu, 3. / 4.) * pow(c->S[iK] / c->S[iRho], 3. / 2.) / l;
        if (is_mu_t && c->S[iEps] != 0) {
            double nu_t = fabs(C_mu * f_mu * c->S[iK] * c->S[iK] / c->S[iEps]);
            c->mu_t = OMIN(nu_t, c->mu_t);
        }
        if (!is_init) {
            c->A[iK] = c->S[iK] * c->U;
            c->A[iEps] = c->S[iEps] * c->U;
            c->B[iK] = c->S[iK] * c->V;
            c->B[iEps] = c->S[iEps] * c->V;
            c->RX[iK] = (c->mu + c->mu_t / sig_k) * c->dkdx;
            c->RX[iEps] = (c->mu + c->mu_t / sig_eps) * c->depsdx;
            c->RY[iK] = (c->mu + c->mu_t / sig_k) * c->dkdy;
            c->RY[iEps] = (c->mu + c->mu_t / sig_eps) * c->depsdy;
            c->A[iK] = c->A[iK] - c->RX[iK];
            c->A[iEps] = c->A[iEps] - c->RX[iEps];
            c->B[iK] = c->B[iK] - c->RY[iK];
            c->B[iEps] = c->B[iEps] - c->RY[iEps];
            if (c->S[iK] != 0.) {
                c->SrcAdd[iK] = c->SrcAdd[iEps] = 0.;
                if (!is_tt(c, TCT_k_CONST))
                    c->Src[iK] = (G - c->S[iEps] * (1 + Mt) + L_k * c->S[iRho]);
                if (!is_tt(c, TCT_eps_CONST) && c->S[iK] != 0)
                    c->Src[iEps] = (C1eps * f1 * c->S[iEps] / c->S[iK] * G -
                                    C2eps * f2 * c->S[iEps] * c->S[iEps] / c->S[iK] + L_eps * c->S[iRho]);
            }
            turb_axisymm_addon(c, g, is_init);
        }
    } else if (is_tt(c, TCT_SA_Model)) {
        double Cb1, Cb2, sig, _k, Cw1, Cw2, Cw3, Cv1, Ct2, Ct4, C5, nu, nu_hat, ksi, fv1, fv2, ft2, fw, gg, r, Omega,
            S_hat, Wxy, Div_nu, a_sound2;
        fv1 = 1.0;
        if (is_init) {
            c->S[iNut] = c->mu / c->S[iRho] / 100.0;
        } else if (is_ct(c, CT_WALL_NO_SLIP) || is_ct(c, CT_WALL_LAW) || is_tt(c, TCT_k_CONST)) {
            c->S[iNut] = 0.;
        } else if (is_ct(c, NT_FC)) {
            c->S[iNut] = c->mu / c->S[iRho] * g->turb_intensity;
        } else {
            Cb1 = 0.1355; Cb2 = 0.622; sig = 2.0 / 3.0; _k = 0.41;
            Cw1 = Cb1 / (_k * _k) + (1 + Cb2) / sig;
            Cw2 = 0.3; Cw3 = 2.0; Cv1 = 7.1; Ct2 = 2.0; Ct4 = 0.5; C5 = 3.5;
            a_sound2 = c->k * c->R * c->Tg;
            nu = c->mu / c->S[iRho];
            ksi = c->S[iNut] / nu;
            fv1 = ksi * ksi * ksi / (ksi * ksi * ksi + Cv1 * Cv1 * Cv1);
            nu_hat = c->mu_t / c->S[iRho] / fv1;
            fv2 = 1.0 - ksi / (1.0 + ksi * fv1);
            Wxy = 0.5 * (c->dVdx - c->dUdy);
            Omega = sqrt(2.0 * Wxy * Wxy);
            S_hat = Omega + c->S[iNut] / (_k * _k * c->l_min * c->l_min) * fv2;
            if (S_hat < 0.3 * Omega) S_hat = 0.3 * Omega;
            r = OMIN((c->S[iNut] / (S_hat * _k * _k * c->l_min * c->l_min)), 10.0);
            gg = r + Cw2 * (pow(r, 6.0) - r);
            fw = gg * pow((1.0 + pow(Cw3, 6.0)) / (pow(gg, 6.0) + pow(Cw3, 6.0)), 1.0 / 6.0);
            ft2 = Ct2 * exp(-Ct4 * ksi * ksi);
            c->A[iNut] = c->S[iNut] * c->U;
            c->B[iNut] = c->S[iNut] * c->V;
            Div_nu = (c->dkdx + c->dkdy);
            c->RX[iNut] = ((c->mu / c->S[iRho] + c->S[iNut]) * c->dkdx) / sig;
            c->RY[iNut] = ((c->mu / c->S[iRho] + c->S[iNut]) * c->dkdy) / sig;
            c->A[iNut] = c->A[iNut] - c->RX[iNut];
            c->B[iNut] = c->B[iNut] - c->RY[iNut];
            c->Src[iNut] = Cb1 * (1.0 - ft2) * S_hat * c->S[iNut] -
                           (Cw1 * fw - Cb1 / (_k * _k) * ft2) * (c->S[iNut] / c->l_min) * (c->S[iNut] / c->l_min) +
                           (Cb2 * Div_nu * Div_nu) / sig - C5 * nu_hat * nu_hat / a_sound2 * c->dUdy * c->dVdx;
        }
        turb_axisymm_addon(c, g, is_init);
        if (is_mu_t) {
            c->mu_t = OMAX(0, (c->S[iRho] * c->S[iNut] * fv1));
            c->lam_t = c->mu_t * c->CP;
        }
    } else if (is_tt(c, TCT_SST_Model)) {
        /* empty in the reference, hyper_flow_node.hpp:919-920 */
    } else if (is_tt(c, TCT_Integral_Model) && c->mu != 0.0) {
        c->Re_local = c->l_min * sqrt(c->S[iRhoU] * c->S[iRhoU] + c->S[iRhoV] * c->S[iRhoV] + 1.e-30) / c->mu;
    } else if (is_tt(c, TCT_Smagorinsky_Model)) {
        double Omega, Wxy, Cs = 0.1;
        double _delta = sqrt(dx * dy);
        Wxy = 0.5 * (c->dVdx - c->dUdy);
        Omega = sqrt(2.0 * Wxy * Wxy);
        if (is_mu_t) {
            c->mu_t = OMAX(0, (c->S[iRho] * (Cs * _delta) * (Cs * _delta) * Omega));
            c->lam_t = c->mu_t * c->CP;
        }
    }
}

/* FlowNode2D::FillNode2D, hyper_flow_node.hpp:374-600 */
static void fill_node(ohf_node* c, const ohf_cfg* g, int is_mu_t, int is_init, double sig_w, double sig_f, int tem,
                      double delta, int sm, int is_src_add) {
    const int a = OHF_NCOMP;
    const double FT = (double)g->flow_type;
    const double dx = g->dx, dy = g->dy;
    const double* Hu = g->Hu;
    int i;
    double sxx, txy, syy, qx, qy, L, _mu, _lam, t00, Tmp1, Tmp2, Tmp3 = 0.;
    _mu = _lam = Tmp2 = 0.;
    if (is_ct(c, CT_SOLID)) return;
    if (c->S[iRho] == 0) return;
    else if (c->k < 1) return;

    c->k = c->CP / (c->CP - c->R);
    if (is_ct(c, CT_U_CONST)) c->S[iRhoU] = c->U * c->S[iRho];
    else c->U = c->S[iRhoU] / c->S[iRho];
    if (is_ct(c, CT_V_CONST)) c->S[iRhoV] = c->V * c->S[iRho];
    else c->V = c->S[iRhoV] / c->S[iRho];

    if (sm == 1) {
        _mu = _lam = 0.;
        if (is_init && c->TurbType > 0) {
            c->mu_t = 5.0 * c->mu;
        } else if (is_init) {
            c->mu_t = c->lam_t = 0.;
        }
        if (c->TurbType > 0) turb_mod_rans(c, g, is_mu_t, is_init, tem, delta);
    }

    Tmp1 = c->S[iRho];
    for (i = 0; i < a; i++) {
        Tmp3 += Hu[i] * c->S[i + 4];
        Tmp1 -= c->S[i + 4];
    }
    Tmp3 += Hu[a] * Tmp1;

    if (is_ct(c, CT_WALL_LAW)) {
        Tmp1 = sqrt(c->U * c->U + c->V * c->V + 1.e-30);
        c->S[iRhoU] = Tmp1 * c->BGX;
        c->S[iRhoV] = Tmp1 * c->BGY;
        c->U = c->S[iRhoU] / c->S[iRho];
        c->V = c->S[iRhoV] / c->S[iRho];
    } else if (is_ct(c, CT_WALL_NO_SLIP)) {
        c->U = c->S[iRhoU] / c->S[iRho];
        c->V = c->S[iRhoV] / c->S[iRho];
        if (is_src_add) {
            c->SrcAdd[iRho] = c->BGX * (c->U - c->Uw) * c->S[iRho] / dx + c->BGY * (c->V - c->Vw) * c->S[iRho] / dy;
            c->SrcAdd[iRhoU] = c->BGX * (c->U - c->Uw) * c->S[iRho];
            c->SrcAdd[iRhoV] = c->BGY * (c->V - c->Vw) * c->S[iRho];
            c->SrcAdd[iRhoE] = 0.;
        } else {
            c->SrcAdd[iRhoU] = c->SrcAdd[iRhoV] = c->SrcAdd[iRhoE] = 0.;
        }
        c->U = c->Uw;
        c->V = c->Vw;
        for (i = 0; i < a; i++) {
            if (is_src_add) c->SrcAdd[4 + i] = c->SrcAdd[iRho] * c->Y[i];
            else c->SrcAdd[4 + i] = 0.;
        }
        c->S[iRhoU] = c->U * c->S[iRho];
        c->S[iRhoV] = c->V * c->S[iRho];
    } else {
        for (i = 0; i < OHF_NEQ; i++) c->SrcAdd[i] = 0.;
    }

    c->p = (c->k - 1.) * (c->S[iRhoE] - c->S[iRho] * (c->U * c->U + c->V * c->V) * 0.5 - Tmp3);
    c->Tg = c->p / c->R / c->S[iRho];

    if (sm == 1) {
        c->lam_t = c->mu_t * c->CP;
        if (is_mu_t) {
            if (is_ct(c, CT_WALL_NO_SLIP) || is_ct(c, CT_WALL_LAW)) {
                _mu = OMAX(0, (c->mu + c->mu_t * sig_w));
                _lam = OMAX(0, (c->lam + c->lam_t * sig_w));
            } else {
                _mu = OMAX(0, (c->mu + c->mu_t * sig_f));
                _lam = OMAX(0, (c->lam + c->lam_t * sig_f));
            }
        } else {
            _mu = c->mu;
            _lam = c->lam;
        }
        c->Diff = _lam / c->CP;
        L = (2. / 3.) * _mu;
        if (g->flow_type == 1) Tmp2 = L * (c->dUdx + c->dVdy + FT * c->V / c->y);
        else Tmp2 = L * (c->dUdx + c->dVdy);
    }

    c->A[iRho] = c->S[iRhoU];
    c->A[iRhoU] = c->p + c->S[iRhoU] * c->U;
    c->A[iRhoV] = c->S[iRhoV] * c->U;
    c->A[iRhoE] = (c->S[iRhoE] + c->p) * c->U;
    c->B[iRho] = c->S[iRhoV];
    c->B[iRhoU] = c->A[iRhoV];
    c->B[iRhoV] = c->p + c->S[iRhoV] * c->V;
    c->B[iRhoE] = (c->S[iRhoE] + c->p) * c->V;
    for (i = 4; i < 4 + a; i++) {
        c->B[i] = c->S[i] * c->V;
        c->A[i] = c->S[i] * c->U;
    }
    if (g->flow_type == 1) {
        c->F[iRho] = FT * c->B[iRho];
        c->F[iRhoU] = FT * c->A[iRhoV];
        c->F[iRhoV] = FT * c->F[iRho] * c->V;
        c->F[iRhoE] = FT * c->B[iRhoE];
        for (i = 4; i < 4 + a; i++) c->F[i] = FT * c->B[i];
    }
    if (sm == 1) {
        sxx = 2. * _mu * c->dUdx - Tmp2;
        syy = 2. * _mu * c->dVdy - Tmp2;
        txy = _mu * (c->dUdy + c->dVdx);
        qx = _lam * c->dTdx;
        qy = _lam * c->dTdy;
        for (i = 0; i < a + 1; i++) {
            qx += c->Diff * (c->CP * c->Tg + Hu[i]) * c->droYdx[i];
            qy += c->Diff * (c->CP * c->Tg + Hu[i]) * c->droYdy[i];
        }
        c->RX[iRho] = 0.;
        c->RX[iRhoU] = sxx;
        c->RX[iRhoV] = txy;
        c->RX[iRhoE] = c->U * sxx + c->V * txy + qx;
        c->RY[iRho] = 0;
        c->RY[iRhoU] = txy;
        c->RY[iRhoV] = syy;
        c->RY[iRhoE] = c->U * txy + c->V * syy + qy;
        c->A[iRhoU] = c->A[iRhoU] - c->RX[iRhoU];
        c->A[iRhoV] = c->A[iRhoV] - c->RX[iRhoV];
        c->A[iRhoE] = c->A[iRhoE] - c->RX[iRhoE];
        c->B[iRhoU] = c->B[iRhoU] - c->RY[iRhoU];
        c->B[iRhoV] = c->B[iRhoV] - c->RY[iRhoV];
        c->B[iRhoE] = c->B[iRhoE] - c->RY[iRhoE];
        for (i = 4; i < 4 + a; i++) {
            c->RX[i] = c->Diff * c->droYdx[i - 4];
            c->RY[i] = c->Diff * c->droYdy[i - 4];
            c->A[i] = c->A[i] - c->RX[i];
            c->B[i] = c->B[i] - c->RY[i];
        }
        if (g->flow_type == 1) {
            t00 = 2 * _mu * c->V / c->y - Tmp2;
            c->F[iRhoU] -= c->RY[iRhoU];
            c->F[iRhoV] -= c->RY[iRhoV] + t00;
            c->F[iRhoE] -= c->RY[iRhoE];
            for (i = 4; i < 4 + a; i++) c->F[i] -= c->RY[i];
        } else {
            for (i = 0; i < OHF_NEQ; i++) c->F[i] = 0.;
        }
    }
}

/* CalcHeatOnWallSources, deeps2d_core.cpp:2679-2833 (single sub-domain: rank==last_rank==0).
 * FlowNode2D::UpNode/DownNode/LeftNode/RightNode are never assigned (always NULL after the
 * constructor's memset, hyper_flow_node.hpp:154-157,285), so num_near_nodes is always 1.   */
static void heat_on_wall_sources(ohf_node* J, const ohf_cfg* g, double dt) {
    const int nx = g->nx, ny = g->ny;
    const double dx = g->dx, dy = g->dy;
    int i, j;
    for (i = 0; i < nx; i++)
        for (j = 0; j < ny; j++) J[(size_t)i * ny + j].Q_conv = 0.;
    for (i = 0; i < nx; i++)
        for (j = 0; j < ny; j++) {
            ohf_node* c = &J[(size_t)i * ny + j];
            ohf_node *up, *down, *left, *right;
            if (is_ct(c, CT_SOLID)) continue;
            up = (j < ny - 1) ? &J[(size_t)i * ny + j + 1] : 0;
            down = (j > 0) ? &J[(size_t)i * ny + j - 1] : 0;
            left = (i > 0) ? &J[(size_t)(i - 1) * ny + j] : 0;
            right = (i < nx - 1) ? &J[(size_t)(i + 1) * ny + j] : 0;
            if (is_ct(c, CT_WALL_LAW) || is_ct(c, CT_WALL_NO_SLIP)) {
                double lam_eff;
                if (down && is_ct(down, CT_SOLID)) {
                    lam_eff = c->lam + c->lam_t;
                    lam_eff = lam_eff / 1;
                    if (down->Q_conv > 0.) down->Q_conv = (down->Q_conv - lam_eff * (down->Tg - c->Tg) / dy) * 0.5;
                    else down->Q_conv = -lam_eff * (down->Tg - c->Tg) / dy;
                    c->SrcAdd[iRhoE] = -dt * down->Q_conv / dy;
                }
                if (up && is_ct(up, CT_SOLID)) {
                    lam_eff = c->lam + c->lam_t;
                    lam_eff = lam_eff / 1;
                    if (up->Q_conv > 0.) up->Q_conv = (up->Q_conv - lam_eff * (up->Tg - c->Tg) / dy) * 0.5;
                    else up->Q_conv = -lam_eff * (up->Tg - c->Tg) / dy;
                    c->SrcAdd[iRhoE] = -dt * up->Q_conv / dy;
                }
                if (left && is_ct(left, CT_SOLID)) {
                    lam_eff = c->lam + c->lam_t;
                    lam_eff = lam_eff / 1;
                    if (left->Q_conv > 0.) left->Q_conv = (left->Q_conv - lam_eff * (left->Tg - c->Tg) / dx) * 0.5;
                    else left->Q_conv = -lam_eff * (left->Tg - c->Tg) / dx;
                    c->SrcAdd[iRhoE] = -dt * left->Q_conv / dx;
                }
                if (right && is_ct(right, CT_SOLID)) {
                    lam_eff = c->lam + c->lam_t;
                    lam_eff = lam_eff / 1;
                    if (right->Q_conv > 0.) right->Q_conv = (right->Q_conv - lam_eff * (right->Tg - c->Tg) / dx) * 0.5;
                    else right->Q_conv = -lam_eff * (right->Tg - c->Tg) / dx;
                    c->SrcAdd[iRhoE] = -dt * right->Q_conv / dx;
                }
            }
        }
}

static int node_active(const ohf_node* c) { /* deeps2d_core.cpp:861-863 */
    return is_ct(c, CT_NODE_IS_SET) && !is_ct(c, CT_SOLID) && !is_ct(c, NT_FC);
}

int ohf_step(ohf_node* J, const ohf_cfg* g, unsigned global_iter, int is_src_add, double global_time, double* dt_min,
             ohf_step_report* rep) {
    const int nx = g->nx, ny = g->ny;
    const double dx = g->dx, dy = g->dy;
    const int NS = (g->problem_type == 1);
    const double dyy = dx / (dx + dy), dxx = dy / (dx + dy);   /* :643-644 */
    const double dx_1 = 1.0 / dx, dy_1 = 1.0 / dy;             /* :720-721 */
    const double beta_scn = ohf_table_getval(&g->tab[OHF_T_BETA], (double)global_iter); /* :737 */
    const double CFL_scn = ohf_table_getval(&g->tab[OHF_T_CFL], (double)global_iter);   /* :738 */
    double RMS[OHF_NEQ], sumDiv[OHF_NEQ], DD_max[OHF_NEQ], sum_RMS[OHF_NEQ], sum_RMS_Div[OHF_NEQ], DD_local[OHF_NEQ];
    unsigned long iRMS[OHF_NEQ], sum_iRMS[OHF_NEQ];
    double max_RMS, dt, dtdx, dtdy;
    int k_max_RMS = -1, i, j, k, i_c = 0, j_c = 0;

    max_RMS = (g->MonitorIndex < 5) ? 0.5 * g->ExitMonitorValue : 0.;   /* :777-780 */
    for (k = 0; k < OHF_NEQ; k++) {
        RMS[k] = sumDiv[k] = DD_max[k] = sum_RMS[k] = sum_RMS_Div[k] = DD_local[k] = 0.;
        iRMS[k] = sum_iRMS[k] = 0;
    }
    dt = *dt_min;                                              /* :803 */
    dtdx = dt / dx;
    dtdy = dt / dy;
    rep->err = 0;

    /* ---- Stage 1, :853-1026 ---- */
    for (i = 0; i < nx; i++)
        for (j = 0; j < ny; j++) {
            ohf_node* c = &J[(size_t)i * ny + j];
            if (node_active(c)) {
                int n1 = (int)c->idXl, n2 = (int)c->idXr, n3 = (int)c->idYu, n4 = (int)c->idYd;
                ohf_node* up = &J[(size_t)i * ny + (j + n3)];
                ohf_node* down = &J[(size_t)i * ny + (j - n4)];
                ohf_node* right = &J[(size_t)(i + n2) * ny + j];
                ohf_node* left = &J[(size_t)(i - n1) * ny + j];
                double n_n = OMAX(n1 + n2, 1), m_m = OMAX(n3 + n4, 1);
                double n_n_1 = 1. / n_n, m_m_1 = 1. / m_m;
                int Num_Eq = OHF_NEQ - add_eq(c);
                int turb2 = NS && (is_tt(c, TCT_k_eps_Model) || is_tt(c, TCT_SA_Model));
                c->time = global_time;
                for (k = 0; k < Num_Eq; k++) {
                    uint64_t cf = 0, dxf = 0, dyf = 0, dx2f = 0, dy2f = 0;
                    int c_flag = 0, dx_flag = 0, dy_flag = 0, dx2_flag = 0, dy2_flag = 0;
                    double beta = c->beta[k], _beta = 1. - beta, dXX, dYY;
                    if (k < 4) {
                        cf = CT_Rho_CONST << k; dxf = CT_dRhodx_NULL << k; dyf = CT_dRhody_NULL << k;
                        dx2f = CT_d2Rhodx2_NULL << k; dy2f = CT_d2Rhody2_NULL << k;
                    } else if (k < 4 + OHF_NCOMP) {
                        cf = CT_Y_CONST; dxf = CT_dYdx_NULL; dyf = CT_dYdy_NULL; dx2f = CT_d2Ydx2_NULL; dy2f = CT_d2Ydy2_NULL;
                    } else if (turb2) {
                        /* quirk (:918-930): the eps branch starts from the eps masks AND shifts them by
                         * (k-7)=1 again, so eps tests 0x04,0x10,0x40,0x100,0x400 -- the last one is the
                         * k-eps model bit itself, i.e. dy2_flag is always set for the eps equation.     */
                        if (k == iK) {
                            cf = TCT_k_CONST << (k - 4 - OHF_NCOMP); dxf = TCT_dkdx_NULL << (k - 4 - OHF_NCOMP);
                            dyf = TCT_dkdy_NULL << (k - 4 - OHF_NCOMP); dx2f = TCT_d2kdx2_NULL << (k - 4 - OHF_NCOMP);
                            dy2f = TCT_d2kdy2_NULL << (k - 4 - OHF_NCOMP);
                        } else if (k == iEps) {
                            cf = TCT_eps_CONST << (k - 4 - OHF_NCOMP); dxf = (TCT_dkdx_NULL << 1) << (k - 4 - OHF_NCOMP);
                            dyf = (TCT_dkdy_NULL << 1) << (k - 4 - OHF_NCOMP); dx2f = (TCT_d2kdx2_NULL << 1) << (k - 4 - OHF_NCOMP);
                            dy2f = (TCT_d2kdy2_NULL << 1) << (k - 4 - OHF_NCOMP);
                        }
                    }
                    if (k < 4 + OHF_NCOMP) {
                        c_flag = !is_ct(c, cf); dx_flag = !is_ct(c, dxf); dy_flag = !is_ct(c, dyf);
                        dx2_flag = is_ct(c, dx2f) ? 1 : 0; dy2_flag = is_ct(c, dy2f) ? 1 : 0;
                    } else if (turb2) {
                        c_flag = !is_tt(c, cf); dx_flag = !is_tt(c, dxf); dy_flag = !is_tt(c, dyf);
                        dx2_flag = is_tt(c, dx2f) ? 1 : 0; dy2_flag = is_tt(c, dy2f) ? 1 : 0;
                    }
                    if (c_flag) {
                        if (dx_flag) {
                            dXX = c->dSdx[k] = (right->A[k] - left->A[k]) * n_n_1;
                        } else {
                            c->S[k] = (left->S[k] * n2 + right->S[k] * n1) * n_n_1;
                            dXX = c->dSdx[k] = 0.;
                        }
                        if (dy_flag) {
                            dYY = c->dSdy[k] = (up->B[k] - down->B[k]) * m_m_1;
                        } else {
                            c->S[k] = (up->S[k] * n3 + down->S[k] * n4) * m_m_1;
                            dYY = c->dSdy[k] = 0;
                        }
                        if (dx2_flag) dXX = (left->dSdx[k] + right->dSdx[k]) * 0.5;
                        if (dy2_flag) dYY = (up->dSdy[k] + down->dSdy[k]) * 0.5;
                        if (g->flow_type) {
                            c->SN[k] = c->S[k] * beta + _beta * (dxx * (left->S[k] + right->S[k]) + dyy * (up->S[k] + down->S[k])) * 0.5 -
                                       (dtdx * dXX + dtdy * (dYY + c->F[k] / (j + 1))) + (c->Src[k]) * dt + c->SrcAdd[k];
                        } else {
                            c->SN[k] = c->S[k] * beta + _beta * (dxx * (left->S[k] + right->S[k]) + dyy * (up->S[k] + down->S[k])) * 0.5 -
                                       (dtdx * dXX + dtdy * dYY) + (c->Src[k]) * dt + c->SrcAdd[k];
                        }
                    }
                }
            }
        }

    /* ---- Stage 2, :1028-1334 ---- */
    for (i = 0; i < nx; i++)
        for (j = 0; j < ny; j++) {
            ohf_node* c = &J[(size_t)i * ny + j];
            if (node_active(c)) {
                int n1 = (int)c->idXl, n2 = (int)c->idXr, n3 = (int)c->idYu, n4 = (int)c->idYd;
                ohf_node* up = &J[(size_t)i * ny + (j + n3)];
                ohf_node* down = &J[(size_t)i * ny + (j - n4)];
                ohf_node* right = &J[(size_t)(i + n2) * ny + j];
                ohf_node* left = &J[(size_t)(i - n1) * ny + j];
                double n_n = OMAX(n1 + n2, 1), m_m = OMAX(n3 + n4, 1);
                double dx_1_n_n_1 = dx_1 / n_n, dy_1_m_m_1 = dy_1 / m_m;
                int Num_Eq = OHF_NEQ - add_eq(c);
                int turb2 = NS && (is_tt(c, TCT_k_eps_Model) || is_tt(c, TCT_SA_Model));
                for (k = 0; k < Num_Eq; k++) {
                    uint64_t c_flag = 0;
                    if (k < 4) c_flag = CT_Rho_CONST << k;
                    else if (k < 4 + OHF_NCOMP) c_flag = CT_Y_CONST;
                    else if (turb2) c_flag = TCT_k_CONST << (k - 4 - OHF_NCOMP);
                    /* quirk: the residual gate tests the CT word even for k>=7 (:1075) */
                    if (!is_ct(c, c_flag) && c->S[k] != 0.) {
                        double Tmp = c->S[k], beta_min, sqrt_RES = 0, absDD;
                        absDD = c->SN[k] - c->S[k];
                        if (fabs(Tmp) > 1.e-15) {
                            DD_local[k] = fabs(absDD / Tmp);
                            sqrt_RES = sqrt(DD_local[k]);
                        } else {
                            DD_local[k] = 1.0;
                        }
                        beta_min = OMIN(g->beta0, beta_scn);
                        if (is_ct(c, CT_NONREFLECTED)) beta_min = g->nrbc_beta0;
                        switch (g->bFF) { /* :1104-1121 */
                        case 0: c->beta[k] = OMIN(beta_min, (beta_min * beta_min) / (beta_min + DD_local[k])); break;
                        case 1: c->beta[k] = OMIN((beta_min + c->beta[k]) * 0.5, (beta_min * beta_min) / (beta_min + DD_local[k])); break;
                        case 2: c->beta[k] = OMIN(beta_min, (beta_min * beta_min) / (beta_min + DD_local[k] * DD_local[k])); break;
                        case 3: c->beta[k] = OMIN((beta_min + c->beta[k]) * 0.5, (beta_min * beta_min) / (beta_min + DD_local[k] * DD_local[k])); break;
                        case 4: c->beta[k] = OMIN(beta_min, (beta_min * beta_min) / (beta_min + sqrt_RES)); break;
                        case 5: c->beta[k] = OMIN((beta_min + c->beta[k]) * 0.5, (beta_min * beta_min) / (beta_min + sqrt_RES)); break;
                        default: break;
                        }
                        if (g->isAlternateRMS) { /* :1138-1143 */
                            sumDiv[k] += Tmp * Tmp;
                            RMS[k] += absDD;
                        } else {
                            RMS[k] += DD_local[k] * DD_local[k];
                        }
                        iRMS[k]++;
                        DD_max[k] = OMAX(DD_max[k], DD_local[k]);
                        if (DD_max[k] == DD_local[k]) { i_c = i; j_c = j; }
                    }
                    if (k < 4 + OHF_NCOMP) {                /* commit, :1156-1164 */
                        if (!is_ct(c, c_flag)) c->S[k] = c->SN[k];
                    } else if (turb2) {
                        if (!is_tt(c, c_flag)) c->S[k] = c->SN[k];
                    }
                }
                if (NS) { /* gradients with mixed new (left, down) / old (right, up) neighbours, :1169-1238 */
                    double aR = right->S[iRho], aL = left->S[iRho], aU = up->S[iRho], aD = down->S[iRho];
                    c->droYdx[OHF_NCOMP] = c->droYdy[OHF_NCOMP] = 0.;
                    for (k = 4; k < OHF_NEQ - 2; k++) {
                        if (!is_ct(c, CT_dYdx_NULL)) {
                            c->droYdx[k - 4] = (right->S[k] - left->S[k]) * dx_1_n_n_1;
                            aR -= right->S[k];
                            aL -= left->S[k];
                        }
                        if (!is_ct(c, CT_dYdy_NULL)) {
                            c->droYdy[k - 4] = (up->S[k] - down->S[k]) * dy_1_m_m_1;
                            aU -= up->S[k];
                            aD -= down->S[k];
                        }
                    }
                    if (!is_ct(c, CT_dYdx_NULL)) c->droYdx[OHF_NCOMP] = (aR - aL) * dx_1_n_n_1;
                    if (!is_ct(c, CT_dYdy_NULL)) c->droYdy[OHF_NCOMP] = (aU - aD) * dy_1_m_m_1;
                    if (is_ct(c, CT_WALL_NO_SLIP) || is_ct(c, CT_WALL_LAW)) {
                        c->dUdx = (right->U * n1 - left->U * n2) * dx_1_n_n_1;
                        c->dVdx = (right->V * n1 - left->V * n2) * dx_1_n_n_1;
                        c->dUdy = (up->U * n3 - down->U * n4) * dy_1_m_m_1;
                        c->dVdy = (up->V * n3 - down->V * n4) * dy_1_m_m_1;
                        if (is_tt(c, TCT_k_eps_Model)) {
                            c->dkdx = (right->S[iK] * n1 - left->S[iK] * n2) * dx_1_n_n_1 / c->S[iRho];
                            c->depsdx = (right->S[iEps] * n1 - left->S[iEps] * n2) * dx_1_n_n_1 / c->S[iRho];
                            c->dkdy = (up->S[iK] * n3 - down->S[iK] * n4) * dy_1_m_m_1 / c->S[iRho];
                            c->depsdy = (up->S[iEps] * n3 - down->S[iEps] * n4) * dy_1_m_m_1 / c->S[iRho];
                        } else if (is_tt(c, TCT_SA_Model)) {
                            c->dkdx = (right->S[iK] * n1 - left->S[iK] * n2) * dx_1_n_n_1 / c->S[iRho];
                            c->dkdy = (up->S[iK] * n3 - down->S[iK] * n4) * dy_1_m_m_1 / c->S[iRho];
                        }
                    } else {
                        c->dUdx = (right->U - left->U) * dx_1_n_n_1;
                        c->dVdx = (right->V - left->V) * dx_1_n_n_1;
                        c->dUdy = (up->U - down->U) * dy_1_m_m_1;
                        c->dVdy = (up->V - down->V) * dy_1_m_m_1;
                        if (is_tt(c, TCT_k_eps_Model)) {
                            c->dkdx = (right->S[iK] - left->S[iK]) * dx_1_n_n_1 / c->S[iRho];
                            c->depsdx = (right->S[iEps] - left->S[iEps]) * dx_1_n_n_1 / c->S[iRho];
                            c->dkdy = (up->S[iK] - down->S[iK]) * dy_1_m_m_1 / c->S[iRho];
                            c->depsdy = (up->S[iEps] - down->S[iEps]) * dy_1_m_m_1 / c->S[iRho];
                        } else if (is_tt(c, TCT_SA_Model)) {
                            c->dkdx = (right->S[iK] - left->S[iK]) * dx_1_n_n_1 / c->S[iRho];
                            c->dkdy = (up->S[iK] - down->S[iK]) * dy_1_m_m_1 / c->S[iRho];
                        }
                    }
                    c->dTdx = (right->Tg - left->Tg) * dx_1_n_n_1;
                    c->dTdy = (up->Tg - down->Tg) * dy_1_m_m_1;
                }
                if ((int)global_iter < g->TurbStartIter) /* :1240-1244 */
                    fill_node(c, g, 0, g->isTurbulenceReset, g->SigW, g->SigF, g->TurbExtModel, g->delta_bl, g->problem_type, is_src_add);
                else
                    fill_node(c, g, 1, 0, g->SigW, g->SigF, g->TurbExtModel, g->delta_bl, g->problem_type, is_src_add);
                if (c->Tg < 0.) { /* :1246-1316: the reference aborts here */
                    rep->err = 1; rep->err_i = i; rep->err_j = j;
                    return 1;
                } else {
                    double CFL_min = OMIN(g->CFL, CFL_scn);
                    double AAA = sqrt(c->k * c->R * c->Tg);
                    double dt_min_local = CFL_min * OMIN(dx / (AAA + fabs(c->U)), dy / (AAA + fabs(c->V)));
                    *dt_min = OMIN(*dt_min, dt_min_local);
                    chemistry(c, g);
                }
            } else if (is_ct(c, NT_FC)) {
                fill_node(c, g, 1, 0, g->SigW, g->SigF, g->TurbExtModel, g->delta_bl, g->problem_type, is_src_add);
            }
        }

    if (!g->isAdiabaticWall) heat_on_wall_sources(J, g, dt); /* :1402 */

    /* residual norms, single sub-domain, :1539-1574 */
    for (k = 0; k < OHF_NEQ; k++) {
        if (g->isAlternateRMS) {
            if (sumDiv[k] > 0.0) { sum_RMS[k] += RMS[k]; sum_RMS_Div[k] += sumDiv[k]; }
            if (sum_RMS_Div[k] > 0.0 && sum_RMS[k] > 0.0) sum_RMS[k] = sqrt(sum_RMS[k] / sum_RMS_Div[k]);
            else sum_RMS[k] = 0.0;
        } else {
            if (iRMS[k] > 0) { sum_RMS[k] += RMS[k]; sum_iRMS[k] += iRMS[k]; }
            if (sum_iRMS[k] != 0 && sum_RMS[k] > 0.0) sum_RMS[k] = sqrt(sum_RMS[k] / sum_iRMS[k]);
            else sum_RMS[k] = 0;
        }
        if (g->MonitorIndex == 0 || g->MonitorIndex > 4) {
            max_RMS = OMAX(sum_RMS[k], max_RMS);
            if (max_RMS == sum_RMS[k]) k_max_RMS = k;
        } else {
            max_RMS = OMAX(sum_RMS[g->MonitorIndex - 1], max_RMS);
            if (max_RMS == sum_RMS[g->MonitorIndex - 1]) k_max_RMS = g->MonitorIndex - 1;
        }
    }
    rep->dt = dt;
    rep->dt_min = *dt_min;
    for (k = 0; k < OHF_NEQ; k++) { rep->sum_RMS[k] = sum_RMS[k]; rep->DD_max[k] = DD_max[k]; }
    rep->max_RMS = max_RMS;
    rep->k_max_RMS = k_max_RMS;
    rep->i_c = i_c;
    rep->j_c = j_c;
    return 0;
}

int ohf_run_cycle(ohf_node* J, const ohf_cfg* g, unsigned last_iter, int nsteps, double global_time, double* dt_min,
                  ohf_step_report* rep_last, double* time_part) {
    int it, is_src_add = (g->AddSrcStartIter < 0 + last_iter) ? 1 : 0; /* :727-731, evaluated with the stale iter... */
    double tp = 0.;
    for (it = 0; it < nsteps; it++) {
        int rc = ohf_step(J, g, (unsigned)it + last_iter, is_src_add, global_time, dt_min, rep_last);
        if (rc) return rc;
        tp += rep_last->dt;
    }
    if (time_part) *time_part = tp;
    return 0;
}

/* ---- wall nodes and wall distance (init-time callers of the step: SURVEY section 8 f-3) -------------------------------
 * Planar interface (x-major planes, node (i,j) at i*ny + j) so that the checker needs no record layout. */

/* GetWallNodes, deeps2d_core.cpp:2081-2104: rows outermost (j), columns inside (i); non-solid nodes carrying the
 * wall-law or the no-slip bit.  out_ij receives (i,j) pairs; returns the count (out_ij may be NULL to count only). */
long ohf_get_wall_nodes(const uint64_t* ct, int nx, int ny, int* out_ij) {
    long n = 0;
    int i, j;
    for (j = 0; j < ny; j++)
        for (i = 0; i < nx; i++) {
            uint64_t c = ct[(long)i * ny + j];
            if ((c & CT_SOLID) == CT_SOLID) continue;
            if ((c & CT_WALL_LAW) == CT_WALL_LAW || (c & CT_WALL_NO_SLIP) == CT_WALL_NO_SLIP) {
                if (out_ij) { out_ij[2 * n] = i; out_ij[2 * n + 1] = j; }
                n++;
            }
        }
    return n;
}

/* SetMinDistanceToWall2D, deeps2d_core.cpp:4783-4832.  Every set, non-solid node: a node with Tg != 0 and p == 0 is
 * turned into a solid one (status 2, CT updated in place as the reference does); every other one (status 1) folds the
 * wall list IN ORDER: l = min(l, d); a wall that attains the running minimum -- ties included, so the LAST of several
 * equidistant walls wins -- becomes (i_wall, j_wall); after every wall l is clamped from below to min(dx, dy).
 * l_min / i_wall / j_wall of all other nodes (status 0) are left as they are. */
void ohf_set_min_distance_to_wall(uint64_t* ct, const double* Tg, const double* p, int nx, int ny, double dx, double dy,
                                  double x0, const int* wall_ij, long n_wall, double* l_min, int* i_wall, int* j_wall,
                                  unsigned char* status) {
    const double min_l_min = OMIN(dx, dy);
    int i, j;
    long ii;
    for (i = 0; i < nx; i++)
        for (j = 0; j < ny; j++) {
            const long n = (long)i * ny + j;
            if (status) status[n] = 0;
            if (!((ct[n] & CT_NODE_IS_SET) == CT_NODE_IS_SET) || (ct[n] & CT_SOLID) == CT_SOLID) continue;
            if (Tg[n] != 0 && p[n] == 0.) {
                ct[n] |= CT_SOLID;
                if (status) status[n] = 2;
                continue;
            }
            if (status) status[n] = 1;
            l_min[n] = OMAX((x0 + dx * (unsigned)nx), (dy * (unsigned)ny));
            {
                const double x = x0 + i * dx, y = j * dy;
                for (ii = 0; ii < n_wall; ii++) {
                    const unsigned iw = (unsigned)wall_ij[2 * ii], jw = (unsigned)wall_ij[2 * ii + 1];
                    const double wx = iw * dx, wy = jw * dy;
                    const double _l_min = sqrt((x - wx) * (x - wx) + (y - wy) * (y - wy));
                    l_min[n] = OMIN(l_min[n], _l_min);
                    if (l_min[n] == _l_min) { i_wall[n] = (int)iw; j_wall[n] = (int)jw; }
                    l_min[n] = OMAX(min_l_min, l_min[n]);
                }
            }
        }
}

/* Recalc_y_plus, deeps2d_core.cpp:2364-2386 (the serial build's variant; start-up, hf2d_start.cpp:299, and once per
 * output cycle on rank 0 of the MPI build, :1723-1726): y+ of every set, non-solid node from the friction velocity of
 * ITS nearest wall node (i_wall, j_wall) and -- as the reference has it -- the grid step min(dx, dy), not l_min. */
void ohf_recalc_y_plus(const uint64_t* ct, int nx, int ny, double dx, double dy, const int* i_wall, const int* j_wall,
                       const double* dUdy, const double* dVdx, const double* mu, const double* rho, double* y_plus) {
    int i, j;
    for (i = 0; i < nx; i++)
        for (j = 0; j < ny; j++) {
            const long n = (long)i * ny + j;
            if ((ct[n] & CT_NODE_IS_SET) == CT_NODE_IS_SET && !((ct[n] & CT_SOLID) == CT_SOLID)) {
                const long w = (long)(unsigned)i_wall[n] * ny + (unsigned)j_wall[n];
                const double tau_w = (fabs(dUdy[w]) + fabs(dVdx[w])) * mu[w];
                if (rho[w] > 0.0 && tau_w > 0.0) {
                    const double U_w = sqrt(tau_w / rho[w] + 1e-30);
                    y_plus[n] = fabs(U_w * OMIN(dx, dy) * rho[n] / mu[n]);
                } else {
                    y_plus[n] = 0.0;
                }
            }
        }
}
