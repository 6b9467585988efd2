// Compile-time shape of one OCP and the kernel's internal variable ordering.
//
// Transcription sizes follow polympc's ContinuousOCP as parameterised by the reference
// (/root/reference/mpc_solver/robot_ocp.hpp:31-42: POLY_ORDER, NUM_SEG, NX=14, NU=7, NP=1, NG=8).
//
// KKT ordering ("kernel order", SURVEY.md A.6 developed further): the reduced KKT matrix
//     S = H + sigma I + diag(rho_box) + A^T diag(rho_A) A
// couples segment s (nodes s*PO .. s*PO+PO) to its neighbours only through the junction states x_{s*PO} and the
// scalar p.  Variables are therefore split into
//     NS interior blocks  I_s = { [x_0 if s=0], u_{sPO}, w_{sPO+1}, .., w_{sPO+PO-1}, [w_{N-1} if s=NS-1] },  w_k=(q,v,u)_k
//     one separator       B   = { x_{PO}, x_{2PO}, .., x_{(NS-1)PO}, p }
// and stored [I_0 | I_1 | .. | I_{NS-1} | B]; p is the last variable (as in polympc's layout).
#pragma once
#include <math.h>
#include "mpcb_common.h"

namespace mpcb {

template <int PO_, int NS_>
struct Shape {
    static constexpr int PO = PO_, NS = NS_;
    static constexpr int NN = PO * NS + 1;        // nodes
    static constexpr int NCOL = NN - 1;           // collocated nodes (final node is free)
    static constexpr int NVAR = 21 * NN + 1;
    static constexpr int NEQ = 14 * NCOL;
    static constexpr int NINEQ = 8 * NN;
    static constexpr int NA = NEQ + NINEQ;
    static constexpr int NDUAL = NA + NVAR;
    static constexpr int NB = 14 * (NS - 1) + 1;  // separator size
    static constexpr int NI = NVAR - NB;          // all interior variables
    static constexpr int MB = 7 + 21 * (PO - 1);  // interior size of a middle segment
    static constexpr int NB0 = (NS > 1 ? 14 : 0) + 1;   // boundary columns of the first / last segment
    static constexpr int NBM = 29;                       // boundary columns of a middle segment
    static constexpr int MMAX = MB + 21 + (NS == 1 ? 14 : 0);
    static constexpr int AUGMAX = MMAX + (NS > 2 ? NBM : NB0);
    static constexpr int VP = ((NA > NVAR ? NA : NVAR) + 7) / 8 * 8;   // padded vector length
    // threads per CTA of the SQP-stage kernels: thread r owns constraint row r and thread i owns variable i, so NT >= max(NA, NVAR)
    static constexpr int NT = (NA > 512 || NVAR > 512) ? 1024 : 512;
    static_assert(NA <= 1024 && NVAR <= 1024, "one thread per constraint row / variable: at most 1024 of each");

#if defined(__CUDACC__)
    __host__ __device__
#endif
    static constexpr int tri(int n) { return n * (n + 1) / 2; }

    MPCB_HD static int m(int s) { return MB + (s == 0 ? 14 : 0) + (s == NS - 1 ? 21 : 0); }
    MPCB_HD static int nb(int s) { return (s > 0 ? 14 : 0) + (s < NS - 1 ? 14 : 0) + 1; }
    MPCB_HD static int ioff(int s) { return s * MB + (s > 0 ? 14 : 0); }
    MPCB_HD static int woff(int s) { return (s > 0 ? tri(MB + 14) : 0) + (s > 1 ? (s - 1) * tri(MB) : 0); }
    MPCB_HD static int foff(int s) { return (s > 0 ? (MB + 14) * NB0 : 0) + (s > 1 ? (s - 1) * MB * NBM : 0); }
    MPCB_HD static int bboff(int s) { return (s > 0 ? tri(NB0) : 0) + (s > 1 ? (s - 1) * tri(NBM) : 0); }
    MPCB_HD static int poff(int s) { return (s > 0 ? NB0 : 0) + (s > 1 ? (s - 1) * NBM : 0); }

    static constexpr int WSIZE = (NS == 1) ? tri(MB + 35) : tri(MB + 14) + (NS - 2) * tri(MB) + tri(MB + 21);
    static constexpr int FSIZE = (NS == 1) ? (MB + 35) : (MB + 14) * NB0 + (NS - 2) * MB * NBM + (MB + 21) * NB0;
    static constexpr int CSIZE = tri(NB);
    static constexpr int BBSIZE = (NS == 1) ? 1 : 2 * tri(NB0) + (NS - 2) * tri(NBM);
    static constexpr int PARTSIZE = (NS == 1) ? 1 : 2 * NB0 + (NS - 2) * NBM;
    static constexpr int MATSIZE = WSIZE + FSIZE + CSIZE;
    // full (square, coalesced) copies of the operators for the shape-generic ADMM iteration (mpcb_gadmm.h): W_s as m x m, F_s
    // column-major (m x nb), C as NB x NB
    MPCB_HD static int wfoff(int s) { int o = 0; for (int t = 0; t < s; t++) o += m(t) * m(t); return o; }
    static constexpr int WFSIZE = (NS == 1) ? (MB + 35) * (MB + 35) : (MB + 14) * (MB + 14) + (NS - 2) * MB * MB + (MB + 21) * (MB + 21);
    static constexpr int FULLSIZE = WFSIZE + FSIZE + NB * NB;

    // separator index (0..NB-1) of boundary column `col` of segment s
    MPCB_HD static int bsep(int s, int col) {
        const int nfirst = (s > 0 ? 14 : 0);
        if (col < nfirst) return 14 * (s - 1) + col;
        if (col < nb(s) - 1) return 14 * s + (col - nfirst);
        return NB - 1;
    }

    // ---- augmented (segment-local) index of a variable of node k seen from segment s:  [0,m) interior, m+col boundary.
    // comp in 0..20: 0-13 state (q,v), 14-20 control.  Node k must belong to segment s (k in [sPO, sPO+PO]).
    MPCB_HD static int aug_index(int s, int k, int comp) {
        const int j = k - s * PO;
        const int base = (s == 0 ? 14 : 0);
        if (j == 0) {
            if (comp >= 14) return base + (comp - 14);                 // u_{sPO} is interior
            return (s == 0) ? comp : m(s) + comp;                     // x_0 interior; x_{sPO} first boundary block
        }
        if (j < PO) return base + 7 + 21 * (j - 1) + comp;
        // j == PO : end node of the segment
        if (s == NS - 1) return m(s) - 21 + comp;                     // w_{N-1} lives in the last interior
        return m(s) + (s > 0 ? 14 : 0) + comp;                        // x_{(s+1)PO}: second boundary block (state only)
    }
    // kernel-order index of x_{k,c} (c in 0..13) / u_{k,j}
    MPCB_HD static int kvar(int k, int comp) {
        if (k == NN - 1) return ioff(NS - 1) + m(NS - 1) - 21 + comp;
        const int s = k / PO, j = k - s * PO;
        if (j == 0 && s > 0 && comp < 14) return NI + 14 * (s - 1) + comp;
        const int a = aug_index(s, k, comp);
        return ioff(s) + a;
    }
    static constexpr int IP = NVAR - 1;   // p

    // decode a kernel-order index: returns node k and comp (0..20), or k = -1 for p.  seg/local of interior vars.
    MPCB_HD static void decode(int i, int& k, int& comp, int& seg, int& loc) {
        if (i >= NI) {
            seg = -1; loc = i - NI;
            if (i == IP) { k = -1; comp = 0; return; }
            const int s = loc / 14 + 1;
            k = s * PO; comp = loc % 14;
            return;
        }
        int s = 0;
        if (i >= MB + 14) { s = 1 + (i - (MB + 14)) / MB; if (s > NS - 1) s = NS - 1; }
        seg = s; loc = i - ioff(s);
        int l = loc;
        if (s == 0) { if (l < 14) { k = 0; comp = l; return; } l -= 14; }
        if (l < 7) { k = s * PO; comp = 14 + l; return; }
        l -= 7;
        const int j = 1 + l / 21;
        k = (j < PO) ? s * PO + j : NN - 1;
        comp = l % 21;
    }
};

// Chebyshev-Gauss-Lobatto differentiation matrix on ascending nodes and the nodes themselves (polympc
// Chebyshev<PO, GAUSS_LOBATTO>).  Filled on the host by mpcb_fill_cheb().
struct ChebConst {
    double D[8][8];
    double xi[8];
    double ts;          // t_scale = 1 / (2 NS)
};

inline void mpcb_fill_cheb(int po, int ns, ChebConst* cc) {
    const double pi = 3.14159265358979323846;
    for (int i = 0; i < 8; i++) { cc->xi[i] = 0; for (int j = 0; j < 8; j++) cc->D[i][j] = 0; }
    for (int j = 0; j <= po; j++) cc->xi[j] = -cos(pi * j / po);
    if (po == 3) { cc->xi[0] = -1.0; cc->xi[1] = -0.5; cc->xi[2] = 0.5; cc->xi[3] = 1.0; }
    for (int i = 0; i <= po; i++) {
        const double ci = (i == 0 || i == po) ? 2.0 : 1.0;
        double rowsum = 0;
        for (int j = 0; j <= po; j++) {
            if (i == j) continue;
            const double cj = (j == 0 || j == po) ? 2.0 : 1.0;
            const double sgn = ((i + j) % 2) ? -1.0 : 1.0;
            cc->D[i][j] = (ci / cj) * sgn / (cc->xi[i] - cc->xi[j]);
            rowsum += cc->D[i][j];
        }
        cc->D[i][i] = -rowsum;
    }
    if (po == 3) {
        const double d3[4][4] = {{-19.0 / 6, 4, -4.0 / 3, 0.5}, {-1, 1.0 / 3, 1, -1.0 / 3}, {1.0 / 3, -1, -1.0 / 3, 1}, {-0.5, 4.0 / 3, -4, 19.0 / 6}};
        for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) cc->D[i][j] = d3[i][j];
    }
    cc->ts = 1.0 / (2.0 * ns);
}

}  // namespace mpcb
