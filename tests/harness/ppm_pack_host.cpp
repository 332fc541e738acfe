// Test-only host harness: runs the product's packed-pair layer-synchronous FAST PPM machine
// (fv3net_b200/csrc/ppm_pack.cuh, the per-thread code of the fast fused kernel) on the CPU so
// that tests/test_ppm_fast_host.py can compare it with the oracle without a GPU.
#include "../../fv3net_b200/csrc/ppm_pack.cuh"

namespace {
template <int NF>
struct Col {
    b200cs::PpmPack<(NF + 1) / 2> m;
    static constexpr int NP = (NF + 1) / 2, NL = 2 * NP;  // odd NF: the last lane repeats field 0
    const float* pe1; const float* qp[NF]; const float* pe2; float* out[NF];
    int km, kn, k; float pt; bool opened;

    template <int PH>
    void emit(bool last, bool general) {
        if (!opened && (pe2[0] <= m.pe_hi || last)) { m.open_edge(pe2[0]); opened = true; }
        if (opened) {
            while (k < kn && (pe2[k + 1] <= m.pe_hi || last)) {
                if (pe2[k + 1] <= pe1[0]) {  // output layer entirely above the column: mappm.f90:62-64
                    for (int f = 0; f < NF; f++) out[f][k] = qp[f][0];
                    pt = pe2[k + 1];
                    k++;
                    continue;
                }
                b200cs::f2 o2[NP];
                if (last) m.template edge<true>(pt, pe2[k + 1], 1.0f / (pe2[k + 1] - pt), o2, true, m.Q[PH]);
                else m.template edge<false>(pt, pe2[k + 1], 1.0f / (pe2[k + 1] - pt), o2);
                float o[NL];
                for (int j = 0; j < NP; j++) { o[2 * j] = b200cs::f2_lo(o2[j]); o[2 * j + 1] = b200cs::f2_hi(o2[j]); }
                for (int f = 0; f < NF; f++) out[f][k] = (pt <= pe1[0]) ? qp[f][0] : o[f];  // mappm.f90:62-64
                pt = pe2[k + 1];
                k++;
            }
            if (general) m.close_general(pt, m.Q[PH]);
            else m.close(pt);
        }
    }
    template <int PH>
    void layer(int L) {
        bool general = true;  // boundary layers: literal limiter tests, NaN-safe close
        if (L >= 1 && L <= km - 3) {
            b200cs::f2 qn[NP];
            for (int j = 0; j < NP; j++) qn[j] = b200cs::f2_make(qp[2 * j][L + 2], qp[(2 * j + 1) % NF][L + 2]);
            // the kernel's NaN window: literal limiter tests from the step that first sees a NaN
            // level (as q[L+2]) until three steps later, and on layer 1
            bool nan_in = false;
            for (int f = 0; f < NF; f++) nan_in |= (qp[f][L + 2] != qp[f][L + 2]);
            if (nan_in) nanwin = 4;
            general = L == 1 || nanwin > 0;
            m.template step<PH>(pe1[L + 3], qn, general);
            if (nanwin > 0) nanwin--;
        } else if (L == km - 2) {
            m.template step_bottom<PH>(iv);
        } else if (L == km - 1) {
            m.template step_last<PH>();
        }
        emit<PH>(L == km - 1, general);
    }
    int iv, nanwin;
    void run() {
        float p[5], qA[NL], qB[NL], qC[NL], qD[NL];
        for (int j = 0; j < 5; j++) p[j] = pe1[j];
        for (int f = 0; f < NL; f++) { qA[f] = qp[f % NF][0]; qB[f] = qp[f % NF][1]; qC[f] = qp[f % NF][2]; qD[f] = qp[f % NF][3]; }
        m.init(p, qA, qB, qC, qD, iv);
        nanwin = 0;
        for (int f = 0; f < NL; f++) if (qA[f] != qA[f] || qB[f] != qB[f] || qC[f] != qC[f] || qD[f] != qD[f]) nanwin = 5;
        k = 0; pt = pe2[0]; opened = false;
        for (int L = 0; L < km; L++) {
            switch (L & 3) {
                case 0: layer<0>(L); break;
                case 1: layer<1>(L); break;
                case 2: layer<2>(L); break;
                default: layer<3>(L); break;
            }
        }
    }
};

template <int NF>
void run(int ncol, int km, const float* pe1, const float* const* q1, int kn, const float* pe2,
         float* const* q2, int iv) {
    for (int i = 0; i < ncol; i++) {
        Col<NF> c;
        c.km = km; c.kn = kn; c.iv = iv;
        c.pe1 = pe1 + (size_t)i * (km + 1);
        c.pe2 = pe2 + (size_t)i * (kn + 1);
        for (int f = 0; f < NF; f++) { c.qp[f] = q1[f] + (size_t)i * km; c.out[f] = q2[f] + (size_t)i * kn; }
        c.run();
    }
}
}  // namespace

extern "C" int ppm_pack_host(int nf, int ncol, int km, const float* pe1, const float* const* q1,
                              int kn, const float* pe2, float* const* q2, int iv) {
    if (km < 4) return -1;
    switch (nf) {
        case 1: run<1>(ncol, km, pe1, q1, kn, pe2, q2, iv); return 0;
        case 2: run<2>(ncol, km, pe1, q1, kn, pe2, q2, iv); return 0;
        case 4: run<4>(ncol, km, pe1, q1, kn, pe2, q2, iv); return 0;
    }
    return -2;
}
