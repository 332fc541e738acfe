// Test-only host harness: runs the product's streaming PPM column machine
// (fv3net_b200/csrc/ppm_stream.cuh, the code the CUDA kernels execute per thread)
// on the CPU so that tests/test_ppm_stream_host.py can compare it with the oracle
// without a GPU.  Built by the test with g++ -O2 -ffp-contract=off.  Not product code.
#include "../../fv3net_b200/csrc/ppm_stream.cuh"

namespace {
template <int NF>
struct HostAcc {
    const float* pe1;   // [km+1]
    const float* qp[NF]; // [km] each
    float pe(int j) const { return pe1[j - 1]; }
    float q(int f, int j) const { return qp[f][j - 1]; }
};

template <int NF>
void run(int ncol, int km, const float* pe1, const float* const* q1, int kn, const float* pe2,
         float* const* q2, int iv, int kord) {
    for (int i = 0; i < ncol; i++) {
        HostAcc<NF> acc;
        acc.pe1 = pe1 + (size_t)i * (km + 1);
        for (int f = 0; f < NF; f++) acc.qp[f] = q1[f] + (size_t)i * km;
        b200cs::PpmStream<NF> s;
        s.init(acc, km, iv, kord);
        const float* p2 = pe2 + (size_t)i * (kn + 1);
        for (int k = 0; k < kn; k++) {
            float out[NF];
            s.remap(acc, p2[k], p2[k + 1], out);
            for (int f = 0; f < NF; f++) q2[f][(size_t)i * kn + k] = out[f];
        }
    }
}
}  // namespace

extern "C" int ppm_stream_host(int nf, int ncol, int km, const float* pe1, const float* const* q1,
                               int kn, const float* pe2, float* const* q2, int iv, int kord) {
    if (km < 4 || kord > 6) return -1;
    switch (nf) {
        case 1: run<1>(ncol, km, pe1, q1, kn, pe2, q2, iv, kord); return 0;
        case 2: run<2>(ncol, km, pe1, q1, kn, pe2, q2, iv, kord); return 0;
        case 4: run<4>(ncol, km, pe1, q1, kn, pe2, q2, iv, kord); return 0;
    }
    return -2;
}
