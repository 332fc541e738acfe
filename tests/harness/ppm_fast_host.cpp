// Test-only host harness: runs the product's FAST PPM column machine
// (fv3net_b200/csrc/ppm_fast.cuh, the per-thread code of the fast fused kernel) on the CPU so
// that tests/test_ppm_fast_host.py can compare it with the oracle without a GPU.
// Built by the test with g++ -O2 (FMA contraction allowed, like the FAST device build).
#include "../../fv3net_b200/csrc/ppm_fast.cuh"

namespace {
template <int NF>
struct HostAcc {
    const float* pe1;     // [km+1]
    const float* qp[NF];  // [km] each
    int km;
    float qv[NF], pv;
    void fetch(int jq, int jpe) {
        if (jq > km - 1) jq = km - 1;
        if (jpe > km) jpe = km;
        for (int f = 0; f < NF; f++) qv[f] = qp[f][jq];
        pv = pe1[jpe];
    }
    float q(int f) const { return qv[f]; }
    float pe(float) const { return pv; }
    float pe0() const { return pe1[0]; }
    void rewind() {}
};

template <int NF>
void run(int ncol, int km, const float* pe1, const float* const* q1, int kn, const float* pe2,
         float* const* q2, int iv) {
    for (int i = 0; i < ncol; i++) {
        HostAcc<NF> acc;
        acc.km = km;
        acc.pe1 = pe1 + (size_t)i * (km + 1);
        for (int f = 0; f < NF; f++) acc.qp[f] = q1[f] + (size_t)i * km;
        b200cs::PpmFast<NF> s;
        s.init(acc, km, iv, acc.pe1[km]);
        const float* p2 = pe2 + (size_t)i * (kn + 1);
        s.open(acc, p2[0]);
        for (int k = 0; k < kn; k++) {
            float out[NF];
            s.remap(acc, p2[k], p2[k + 1], 1.0f / (p2[k + 1] - p2[k]), out);
            for (int f = 0; f < NF; f++) q2[f][(size_t)i * kn + k] = out[f];
        }
    }
}
}  // namespace

extern "C" int ppm_fast_host(int nf, int ncol, int km, const float* pe1, const float* const* q1,
                             int kn, const float* pe2, float* const* q2, int iv) {
    if (km < 4) return -1;
    switch (nf) {
        case 1: run<1>(ncol, km, pe1, q1, kn, pe2, q2, iv); return 0;
        case 2: run<2>(ncol, km, pe1, q1, kn, pe2, q2, iv); return 0;
        case 4: run<4>(ncol, km, pe1, q1, kn, pe2, q2, iv); return 0;
    }
    return -2;
}
