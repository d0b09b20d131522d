// gs_noalloc.cu -- gsb_field_stats_noalloc: field statistics for callers that must not allocate.
//
// Who this is for.  gsb_field_stats belongs to a handle, and creating a handle calls cudaMalloc.
// Some users of this library cannot afford that:
//   * a host whose device memory is wholly owned by another allocator -- CUDA.jl's memory pool on
//     the Julia side of the drop-in, PyTorch's caching allocator in the Python harness.  A
//     cudaMalloc issued behind that allocator's back competes with it for the same bytes and can
//     fail although the pool still has room (at L = 2048 on one GPU the fields alone take 142 of
//     180 GB);
//   * code in the middle of the time loop (an in-situ analysis hook between iterate! calls):
//     cudaMalloc / cudaFree may synchronise the whole device, which would stall the halo exchange
//     that gsb_step overlaps with the interior update.
// Such a caller already holds the field as a device array (a CuArray / torch tensor in the
// reference layout) and a place for the five results; all this entry point adds is the scratch
// for the per-plane partial results, and that scratch lives in the library image:
//
//   g_noalloc_scratch : file-scope `static __device__`, GSB_NOALLOC_MAX_PLANES x 5 doubles
//                       (160 KiB).  Every GPU that loads libgsb200.so holds it until exit.
//
// Rules that follow from sharing one array:
//   * fields with more than GSB_NOALLOC_MAX_PLANES z-planes are rejected;
//   * the array keeps the previous call's contents.  Nothing in it has to start at zero: pass 1
//     writes all five values of every plane slot that pass 2 reads, before pass 2 reads them
//     (same stream), so no clearing is needed and none is done;
//   * two calls whose kernels overlap on the GPU (different streams, no ordering between them)
//     race on the array and give garbage.  Order them on one stream or with events.
//   * the first call on each device looks up the array's device address
//     (cudaGetSymbolAddress: the name `g_noalloc_scratch` in host code is only a stand-in, not a
//     device pointer) and remembers it.  Make that first call outside any stream capture.
// The entry point calls no cudaMalloc, cudaFree, or library routine that allocates.
#include <cfloat>
#include <cstdio>
#include <mutex>

#include "../../include/gsb200.h"
#include "gs_common.cuh"

namespace {

constexpr int kMaxPlanes = GSB_NOALLOC_MAX_PLANES;
constexpr int kThreads = 256;

static __device__ __align__(256) double g_noalloc_scratch[kMaxPlanes * 5];

// pass 1: one CTA per z-plane of the interior -> 5 partial results in scratch[plane * 5 ..]
template <typename T>
__global__ void __launch_bounds__(kThreads)
plane_stats_kernel(const T* __restrict__ f, int sx, int sy, double thresh, double* __restrict__ scratch) {
    const int z = blockIdx.x;                                   // interior plane
    const long long nx = sx + 2, nxy = nx * (long long)(sy + 2);
    const T* plane = f + nxy * (long long)(z + 1);
    double s = 0.0, s2 = 0.0, mn = DBL_MAX, mx = -DBL_MAX, cnt = 0.0;
    const long long cells = (long long)sx * sy;
    for (long long c = threadIdx.x; c < cells; c += kThreads) {
        const int y = (int)(c / sx), x = (int)(c % sx);
        const double a = (double)plane[(x + 1) + nx * (long long)(y + 1)];
        s += a; s2 += a * a; mn = fmin(mn, a); mx = fmax(mx, a); cnt += a > thresh ? 1.0 : 0.0;
    }
    __shared__ double sh[5][kThreads];
    sh[0][threadIdx.x] = s; sh[1][threadIdx.x] = s2; sh[2][threadIdx.x] = mn; sh[3][threadIdx.x] = mx;
    sh[4][threadIdx.x] = cnt;
    __syncthreads();
    for (int o = kThreads / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) {
            sh[0][threadIdx.x] += sh[0][threadIdx.x + o];
            sh[1][threadIdx.x] += sh[1][threadIdx.x + o];
            sh[2][threadIdx.x] = fmin(sh[2][threadIdx.x], sh[2][threadIdx.x + o]);
            sh[3][threadIdx.x] = fmax(sh[3][threadIdx.x], sh[3][threadIdx.x + o]);
            sh[4][threadIdx.x] += sh[4][threadIdx.x + o];
        }
        __syncthreads();
    }
    if (threadIdx.x < 5) scratch[z * 5 + threadIdx.x] = sh[threadIdx.x][0];   // every slot, every call
}

// pass 2: fold the plane partials in plane order (deterministic) into the caller's out[5]
__global__ void plane_stats_final_kernel(const double* __restrict__ scratch, int nplanes, double* __restrict__ out) {
    const int q = threadIdx.x;
    if (q >= 5) return;
    double a = q == 2 ? DBL_MAX : (q == 3 ? -DBL_MAX : 0.0);
    for (int z = 0; z < nplanes; ++z) {
        const double x = scratch[z * 5 + q];
        a = q == 2 ? fmin(a, x) : q == 3 ? fmax(a, x) : a + x;
    }
    out[q] = a;
}

std::mutex g_mu;
double* g_scratch_ptr[64] = {};     // device address of g_noalloc_scratch, per device ordinal

// the work; a non-zero result leaves the reason in g_msg
int stats_noalloc(int dtype, const int64_t sizes[3], const void* field_dev, double thresh, double* out_dev,
                  void* stream, char* g_msg, size_t g_len) {
    g_msg[0] = 0;
    if (!sizes || !field_dev || !out_dev) {
        snprintf(g_msg, g_len, "gsb_field_stats_noalloc: null argument");
        return 1;
    }
    if (dtype != GSB_F32 && dtype != GSB_F64) {
        snprintf(g_msg, g_len, "gsb_field_stats_noalloc: dtype must be GSB_F32 or GSB_F64");
        return 1;
    }
    for (int d = 0; d < 3; ++d)
        if (sizes[d] < 1 || sizes[d] > (1 << 30)) {
            snprintf(g_msg, g_len, "gsb_field_stats_noalloc: sizes[%d] = %lld out of range", d, (long long)sizes[d]);
            return 1;
        }
    if (sizes[2] > kMaxPlanes) {
        snprintf(g_msg, g_len,
                 "gsb_field_stats_noalloc: %lld z-planes exceed the built-in scratch of GSB_NOALLOC_MAX_PLANES = %d "
                 "planes; use gsb_field_stats (which allocates) for larger fields",
                 (long long)sizes[2], kMaxPlanes);
        return 1;
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        snprintf(g_msg, g_len, "gsb_field_stats_noalloc: cudaGetDevice -> %s (device %d)", cudaGetErrorString(e), dev);
        return 1;
    }
    double* scratch = nullptr;
    {
        // set-up, once per device, with that device current: resolve the symbol to a device address
        std::lock_guard<std::mutex> lock(g_mu);
        if (!g_scratch_ptr[dev]) {
            double* p = nullptr;
            e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
            if (e != cudaSuccess || !p) {
                snprintf(g_msg, g_len, "gsb_field_stats_noalloc: cudaGetSymbolAddress -> %s", cudaGetErrorString(e));
                return 1;
            }
            g_scratch_ptr[dev] = p;
        }
        scratch = g_scratch_ptr[dev];
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int sx = (int)sizes[0], sy = (int)sizes[1], sz = (int)sizes[2];
    if (dtype == GSB_F32)
        plane_stats_kernel<float><<<sz, kThreads, 0, st>>>((const float*)field_dev, sx, sy, thresh, scratch);
    else
        plane_stats_kernel<double><<<sz, kThreads, 0, st>>>((const double*)field_dev, sx, sy, thresh, scratch);
    plane_stats_final_kernel<<<1, 32, 0, st>>>(scratch, sz, out_dev);
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(g_msg, g_len, "gsb_field_stats_noalloc: launch -> %s", cudaGetErrorString(e));
        return 1;
    }
    return 0;
}

}  // namespace

namespace gsb { void set_last_error(const char* msg); }   // gsb200_api.cu: the message gsb_last_error() returns

// On failure the reason goes to errbuf (when the caller passed one) and, like every other entry point
// of this library, to gsb_last_error() of the calling thread.  On success the two kernels are only
// enqueued: out_dev holds the results once `stream` has completed.
extern "C" int gsb_field_stats_noalloc(int dtype, const int64_t sizes[3], const void* field_dev, double thresh,
                                       double* out_dev, void* stream, char* errbuf, size_t errbuf_len) {
    char msg[320];
    const int rc = stats_noalloc(dtype, sizes, field_dev, thresh, out_dev, stream, msg, sizeof msg);
    if (errbuf && errbuf_len) snprintf(errbuf, errbuf_len, "%s", msg);
    if (rc != 0) gsb::set_last_error(msg);
    return rc;
}
