// gs_tune.cu -- kernel-variant micro-benchmark through the C ABI (no Python, no torch).
//   gs_tune L dtype(f32|f64) arith(ref|native) noise(0|1) steps [variant ...]
// variant = kind:ty:stages:zchunk, e.g. tma:8:4:64 l1:4:0:128.  Prints one line per variant:
// ms/step, GLUPS, algorithmic GB/s, and a checksum (sum u, sum v after the timed steps from the
// same start) so variants can be compared bit for bit.
#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/gsb200.h"

#define OK(x) do { if ((x) != 0) { fprintf(stderr, "FAIL %s: %s\n", #x, gsb_last_error()); return 1; } } while (0)

int main(int argc, char** argv) {
    if (argc < 6) { fprintf(stderr, "usage: gs_tune L f32|f64 ref|native noise steps [kind:ty:stages:zchunk ...]\n"); return 2; }
    const long long L = atoll(argv[1]);
    const int dtype = !strcmp(argv[2], "f64") ? GSB_F64 : GSB_F32;
    const int arith = !strcmp(argv[3], "native") ? GSB_ARITH_NATIVE : GSB_ARITH_REFERENCE;
    const double noise = atof(argv[4]) != 0.0 ? 0.1 : 0.0;
    const int steps = atoi(argv[5]);
    std::vector<std::string> variants;
    for (int i = 6; i < argc; ++i) variants.push_back(argv[i]);
    if (variants.empty()) variants = {"tma:8:4:64", "tma:16:4:64", "tma:16:3:64", "tma:8:6:64", "l1:4:0:64", "l1:8:0:64"};

    gsb_config cfg;
    memset(&cfg, 0, sizeof cfg);
    OK(gsb_fill_domain(&cfg, L, 1, 0));
    cfg.dtype = dtype; cfg.arith = arith;
    cfg.Du = 0.2; cfg.Dv = 0.1; cfg.F = 0.02; cfg.k = 0.048; cfg.dt = 1.0; cfg.noise = noise;  // examples/settings-files.toml
    cfg.seed = 0x5EED; cfg.device = 0;
    gsb_t* h = nullptr;
    OK(gsb_create(&h, &cfg));
    void* sv = nullptr;
    OK(gsb_compute_stream(h, &sv));
    cudaStream_t st = (cudaStream_t)sv;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const double cells = (double)L * L * L, bpc = dtype == GSB_F32 ? 16.0 : 32.0;
    printf("# L=%lld %s %s noise=%g steps=%d\n", L, argv[2], argv[3], noise, steps);
    for (const std::string& v : variants) {
        char kind[16]; int ty = 0, sg = 0, zc = 0;
        if (sscanf(v.c_str(), "%15[^:]:%d:%d:%d", kind, &ty, &sg, &zc) < 2) { fprintf(stderr, "bad variant %s\n", v.c_str()); continue; }
        const int kk = !strcmp(kind, "l1") ? GSB_KERNEL_L1 : !strcmp(kind, "tmalo") ? GSB_KERNEL_TMA_LO : !strcmp(kind, "tmahi") ? GSB_KERNEL_TMA_HI : GSB_KERNEL_TMA;
        if (gsb_set_kernel(h, kk, ty, sg, zc)) {
            printf("%-16s skipped: %s\n", v.c_str(), gsb_last_error());
            continue;
        }
        OK(gsb_init_fields(h));
        OK(gsb_step(h, 3));
        OK(gsb_sync(h));
        cudaEventRecord(e0, st);
        OK(gsb_step(h, steps));
        cudaEventRecord(e1, st);
        OK(gsb_sync(h));
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        double su[5], svv[5];
        OK(gsb_field_stats(h, GSB_U, 0.1, su));
        OK(gsb_field_stats(h, GSB_V, 0.1, svv));
        const double per = ms / steps;
        printf("%-16s %8.4f ms/step %8.2f GLUPS %8.1f GB/s  sum_u=%.17g sum_v=%.17g\n", v.c_str(), per,
               cells / per / 1e6, cells * bpc / per / 1e6, su[0], svv[0]);
        fflush(stdout);
    }
    OK(gsb_destroy(h));

    // GS_TUNE_RANKS=n: n weak-scaled subdomains (L^3 each, the reference decomposition) as in-process
    // ranks on this one device; isolates the cost of the interior/shell split and of pack/unpack
    // from the transport.  Ideal time per step = n x the single-subdomain time.
    const char* er = getenv("GS_TUNE_RANKS");
    const int nr = er ? atoi(er) : 0;
    if (nr > 1) {
        std::vector<gsb_t*> hs(nr, nullptr);
        int32_t dims[3];
        OK(gsb_dims_create(nr, dims));
        for (int r = 0; r < nr; ++r) {
            gsb_config c = cfg;
            OK(gsb_fill_domain(&c, L * dims[0], nr, r));
            for (int d = 0; d < 3; ++d) { c.global[d] = L * dims[d]; c.sizes[d] = L; c.offsets[d] = L * c.coords[d]; }
            c.dtype = dtype; c.arith = arith; c.Du = cfg.Du; c.Dv = cfg.Dv; c.F = cfg.F; c.k = cfg.k; c.dt = cfg.dt;
            c.noise = noise; c.seed = cfg.seed;
            int ndev = 1;
            cudaGetDeviceCount(&ndev);
            c.device = getenv("GS_TUNE_SPREAD") ? r % ndev : 0;   // one subdomain per GPU: peer-store transport
            OK(gsb_create(&hs[r], &c));
            OK(gsb_init_fields(hs[r]));
        }
        OK(gsb_comm_init_local(hs.data(), nr));
        void* sv0 = nullptr;
        OK(gsb_compute_stream(hs[nr - 1], &sv0));
        OK(gsb_step_group(hs.data(), nr, 3));
        for (int r = 0; r < nr; ++r) OK(gsb_sync(hs[r]));
        OK(gsb_timers_enable(hs[0], 1));
        // several devices may be involved: wall clock around fully synchronised steps
        auto t0 = std::chrono::steady_clock::now();
        OK(gsb_step_group(hs.data(), nr, steps));
        for (int r = 0; r < nr; ++r) OK(gsb_sync(hs[r]));
        const float ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
        double t[4];
        OK(gsb_timers(hs[0], t));
        printf("ranks=%d (%s): %8.4f ms/step total = %8.4f ms per subdomain; rank0 last step: halo %.3f interior %.3f shell %.3f step %.3f ms\n",
               nr, getenv("GS_TUNE_SPREAD") ? "one per GPU, peer stores" : "all on device 0", ms / steps, ms / steps / nr, t[0], t[1], t[2], t[3]);
        for (int r = 0; r < nr; ++r) OK(gsb_destroy(hs[r]));
    }
    return 0;
}
