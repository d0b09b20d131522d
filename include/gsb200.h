/*
 * gsb200.h -- C ABI of libgsb200.so: the B200 (sm_100a) backend for GrayScott.jl's per-step
 * hot path (halo exchange -> fused 7-point Laplacian + Gray-Scott reaction (+noise) -> swap).
 *
 * This is the drop-in boundary: exactly what a `backend = "B200"` extension of the reference
 * would `ccall` (see INTEGRATION.md and julia/B200Backend.jl).  Each entry point names the
 * reference interface it replaces; paths are relative to the reference repository root.
 *
 * Conventions
 *  - every function returns 0 on success, non-zero on failure; gsb_last_error() then returns a
 *    message (thread-local).  No C++ exception and no abort crosses this boundary
 *    (the reference maps any failure to exit code 1, src/GrayScott.jl:40-48).
 *  - one handle = one subdomain = one GPU (the reference: one MPI rank, one Fields).  Handles
 *    are used from one host thread at a time.  Several handles per process are allowed.
 *  - step/exchange/calculate/swap (and gsb_calculate_arrays, gsb_field_stats_noalloc on the
 *    caller's stream) only ENQUEUE work; they do not synchronise with the host.
 *    gsb_get_* / gsb_sync / gsb_field_stats / gsb_pdf / gsb_timers / gsb_destroy do.
 *  - host arrays use the REFERENCE layout: x fastest, then y, then z
 *    (Julia Array{T,3}(nx, ny, nz), column-major).  Device memory is owned by the library in
 *    its own padded layout and is never exposed.
 *  - there is no CPU fallback: without a CUDA device every call that needs one fails.
 */
#ifndef GSB200_H
#define GSB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSB_VERSION 100

/* settings.precision, src/simulation/communication.jl:27 */
#define GSB_F32 0
#define GSB_F64 1

/* Arithmetic of the cell update.
 * GSB_ARITH_REFERENCE: rounding-for-rounding the expression of calculate!
 *   (src/simulation/Simulation_CPU.jl:96-109 with Julia's promotion: Float32 fields are
 *   updated mostly in Float64).  Bit-identical to the reference CPU solver at noise = 0.
 * GSB_ARITH_NATIVE: Float32 only; same formula in Float32 FMAs (within 1e-5 of the reference). */
#define GSB_ARITH_REFERENCE 0
#define GSB_ARITH_NATIVE 1

/* stencil kernel variants (gsb_set_kernel) */
#define GSB_KERNEL_TMA 0    /* TMA-staged tiles, occupancy chosen per arithmetic */
#define GSB_KERNEL_L1 1     /* no shared memory, y neighbours through L1 */
#define GSB_KERNEL_TMA_LO 2 /* TMA, register cap for the lower occupancy */
#define GSB_KERNEL_TMA_HI 3 /* TMA, register cap for the higher occupancy */

/* neighbour slots, the order of MPICartDomain.proc_neighbors (communication.jl:90-92) */
#define GSB_WEST 0  /* x-1 */
#define GSB_EAST 1  /* x+1 */
#define GSB_DOWN 2  /* y-1 */
#define GSB_UP 3    /* y+1 */
#define GSB_SOUTH 4 /* z-1 */
#define GSB_NORTH 5 /* z+1 */

/* field selectors for the ghosted accessors: Fields.u / .v / .u_temp / .v_temp
 * (src/simulation/Structs.jl:82-93) as they are bound NOW (the swap re-binds them). */
#define GSB_U 0
#define GSB_V 1
#define GSB_U_TEMP 2
#define GSB_V_TEMP 3

typedef struct gsb_handle gsb_t;

/* What the callers of init_fields / iterate! hold: Settings (src/simulation/Structs.jl:4-28)
 * and MPICartDomain (src/simulation/Structs.jl:57-73, filled by init_domain,
 * src/simulation/communication.jl:59-96). */
typedef struct gsb_config {
    int32_t dtype;        /* GSB_F32 / GSB_F64                                        */
    int32_t arith;        /* GSB_ARITH_*                                              */
    int64_t L;            /* settings.L (global cube edge); see `global` below        */
    int64_t global[3];    /* global extent per axis; {0,0,0} means {L,L,L}.  Non-cubic
                             extents are an extension used by the weak-scaling bench. */
    int32_t dims[3];      /* mcd.dims                                                 */
    int32_t coords[3];    /* mcd.coords                                               */
    int64_t sizes[3];     /* mcd.proc_sizes   (sx, sy, sz)                            */
    int64_t offsets[3];   /* mcd.proc_offsets (0-based global)                        */
    int32_t neighbors[6]; /* mcd.proc_neighbors, GSB_WEST..GSB_NORTH; -1 = MPI.PROC_NULL */
    double Du, Dv, F, k, dt, noise; /* Float64 as in Settings; rounded to T per step as
                                       calculate! does (Simulation_CPU.jl:82-87)      */
    uint64_t seed;        /* noise key (the reference's RNG is unseeded)              */
    int32_t device;       /* CUDA device ordinal                                      */
    int32_t rank, nranks; /* MPI.Comm_rank / Comm_size of the Cartesian communicator  */
} gsb_config;

/* ---- lifetime ------------------------------------------------------------------------- */

int gsb_version(void);
const char* gsb_last_error(void);

/* Allocates the four field arrays (u, v, u_temp, v_temp) on cfg->device and the streams.
 * Does NOT fill them: call gsb_init_fields or gsb_set_ghosted.
 * Replaces the allocations of init_fields (Simulation_CPU.jl:23-28, ext/CUDAExt.jl:40-45). */
int gsb_create(gsb_t** out, const gsb_config* cfg);
int gsb_destroy(gsb_t* h);

/* Convenience: MPI.Dims_create(n, [0,0,0]) + Cart_coords/Cart_shift of a non-periodic,
 * non-reordered Cartesian communicator (init_domain, communication.jl:59-96), for callers
 * without MPI (the Python harness).  Fills dims/coords/sizes/offsets/neighbors/rank/nranks/
 * L/global of *cfg; L must divide by every dims[i] (the reference raises InexactError). */
int gsb_dims_create(int nranks, int32_t dims[3]);
int gsb_fill_domain(gsb_config* cfg, int64_t L, int nranks, int rank);

/* ---- the hot path --------------------------------------------------------------------- */

/* init_fields(::Val{:cpu}, ::Val{:plain}, ...)  src/simulation/Simulation_CPU.jl:14-65:
 * u = 1, v = 0, u_temp = v_temp = 0 (ghosts included), then the 13^3 block around the global
 * centre u = 0.25, v = 0.33.  Resets the step counter to 0.  Runs on the device. */
int gsb_init_fields(gsb_t* h);

/* exchange!(fields, mcd)  src/simulation/communication.jl:138-199: ghost faces of u and v from
 * the neighbours' boundary layers.  No-op for faces whose neighbour is -1 (non-periodic), so
 * ghost cells on the global boundary are never written.  Requires gsb_comm_init* if any
 * neighbour is >= 0. */
int gsb_exchange(gsb_t* h);

/* calculate!(::Val{backend}, ::Val{:plain}, fields, settings, mcd)
 * src/simulation/Simulation_CPU.jl:77-113: u_temp, v_temp <- one step from u, v
 * (interior cells only). */
int gsb_calculate(gsb_t* h);

/* fields.u, fields.u_temp = fields.u_temp, fields.u (and v)  src/simulation/public.jl:67-68.
 * Also advances the step counter (the noise counter). */
int gsb_swap(gsb_t* h);

/* iterate!  src/simulation/public.jl:54-71, nsteps times: exchange -> calculate -> swap, with
 * the exchange overlapped with the interior update. */
int gsb_step(gsb_t* h, int nsteps);

/* Same for a group of handles living in one process and connected with gsb_comm_init_local
 * (each on its own or on a shared device): all handles advance in lockstep. */
int gsb_step_group(gsb_t** hs, int n, int nsteps);

/* ---- data access ---------------------------------------------------------------------- */

/* get_fields(::Val{backend}, fields)  src/simulation/Simulation_CPU.jl:125-133 /
 * ext/CUDAExt.jl:199-209: ghost-stripped copies of u and v, sx*sy*sz each, to HOST memory
 * (pinned memory makes it faster; pageable works).  Synchronises. */
int gsb_get_fields(gsb_t* h, void* u_host, void* v_host);

/* The same copies without stalling the time loop (IO.write_step! runs every `plotgap` steps,
 * src/simulation/IO.jl:82-96, src/GrayScott.jl:86-95): a ghost-free snapshot is taken on the device
 * in stream order, then drained to the (preferably pinned) host buffers on a copy stream while
 * later steps run.  The host buffers are valid after gsb_output_wait (or the next
 * gsb_get_fields_async, which waits for the previous one).  Needs 2 * sx*sy*sz * sizeof(T) bytes of
 * extra device memory, allocated on first use; when that does not fit it degrades to gsb_get_fields. */
int gsb_get_fields_async(gsb_t* h, void* u_host, void* v_host);
int gsb_output_wait(gsb_t* h);

/* Array(fields.<which>) / copyto!: the ghosted (sx+2)(sy+2)(sz+2) array in the reference
 * layout, from / to HOST memory.  For tests and restart.  Synchronise. */
int gsb_get_ghosted(gsb_t* h, int which, void* host);
int gsb_set_ghosted(gsb_t* h, int which, const void* host);

/* pinned host memory for the accessors above (cudaHostAlloc / cudaFreeHost) */
int gsb_alloc_host(void** out, size_t bytes);
int gsb_free_host(void* p);

int gsb_sync(gsb_t* h);
/* The step counter (= number of swaps since gsb_init_fields): the noise counter, and the parity that
 * selects receive buffers.  Set it together with gsb_set_ghosted to resume a run (restart,
 * Structs.jl:15-19): the next step then exchanges the ghosts afresh. */
int gsb_get_step(gsb_t* h, int64_t* step);
int gsb_set_step(gsb_t* h, int64_t step);

/* sum, sum of squares, min, max, count(value > thresh) over the interior cells of field
 * `which` (out[0..4]); deterministic.  For the statistical noise checks at full size. */
int gsb_field_stats(gsb_t* h, int which, double thresh, double out[5]);

/* _compute_pdf(data, shape, count, nbins, min, max)  src/analysis/pdfcalc.jl:13-49 (unfinished in the
 * reference: no return, 0-based bin into a 1-based array): for every z slice of field `which` the
 * number of interior cells per bin, nbins equal bins over [min, max], bin = floor((value - min) /
 * bin_width) evaluated in T, value == max in the last bin, values outside the window not counted;
 * nbins == 1 or a window narrower than 1e-20 gives slice_size everywhere, like the reference.
 * pdf_host: nbins * sz counts, bin fastest; bins_host (optional): the nbins lower bin edges. */
int gsb_pdf(gsb_t* h, int which, int nbins, double min, double max, uint32_t* pdf_host, double* bins_host);

/* ---- communication set-up (replaces MPI.Init / Cart_create for the device side) -------- */

/* NCCL transport, one process per GPU: rank 0 obtains the 128-byte id, the caller broadcasts it
 * (MPI.Bcast in Julia, torch.distributed in the Python harness), every rank calls
 * gsb_comm_init.  Neighbour numbers in gsb_config are NCCL ranks. */
int gsb_get_unique_id(void* id128);
int gsb_comm_init(gsb_t* h, const void* id128);

/* In-process transport: hs[r] is the handle of rank r (all nranks of them).  Faces are written
 * directly into the neighbour's receive buffers (same device, or peer device over NVLink). */
int gsb_comm_init_local(gsb_t** hs, int n);

/* ---- tuning / measurement ------------------------------------------------------------- */

/* Select the stencil kernel: kind GSB_KERNEL_*, rows per CTA, TMA ring depth, planes per CTA.
 * Non-positive values keep the current setting.  Also read from $GSB_KERNEL="kind:ty:stages:zchunk"
 * at gsb_create. */
int gsb_set_kernel(gsb_t* h, int kind, int ty, int stages, int zchunk);
int gsb_get_kernel(gsb_t* h, int* kind, int* ty, int* stages, int* zchunk);

/* Per-phase device times of the LAST step enqueued by gsb_step (ms): out[0] = halo (pack,
 * transport, unpack; summed over the pieces of the step, see gsb_halo_timers), out[1] = interior
 * update, out[2] = shell update after the halo arrived (split scheme only; the pipelined scheme
 * has no shell kernels), out[3] = whole step.
 * Enable before stepping; synchronises when read. */
int gsb_timers_enable(gsb_t* h, int on);
int gsb_timers(gsb_t* h, double out_ms[4]);

/* The halo exchange of the LAST step piece by piece (the reference: the 12 MPI.Sendrecv! of
 * communication.jl:146-177; here one piece per z chunk of the pipelined scheme, or one piece for the
 * whole faces): device time of the pack kernel, of the transport (NCCL send/recv group, which
 * includes waiting for the neighbour to reach the same point; 0 for in-process ranks, whose pack
 * kernel writes into the neighbour's buffer directly) and of the unpack kernel, plus the bytes this
 * rank sent with the piece.  Arrays of max_pieces entries (any may be NULL); *npieces = entries
 * filled.  Enable the timers before stepping; synchronises when read. */
int gsb_halo_timers(gsb_t* h, int max_pieces, double* pack_ms, double* wire_ms, double* unpack_ms,
                    int64_t* bytes, int* npieces);

/* cudaGetDeviceCount, for hosts that map node-local ranks to devices without a CUDA binding. */
int gsb_device_count(int* count);

/* Halo bytes this handle sends per step (u and v faces to all real neighbours). */
int gsb_halo_bytes(gsb_t* h, int64_t* bytes);

/* Number of kernels this handle has launched so far (update, pack and unpack kernels; NCCL's own
 * kernels are not counted). */
int gsb_launch_count(gsb_t* h, int64_t* count);

/* Stream on which gsb_step's update kernels run (a cudaStream_t), for callers that time with
 * their own CUDA events. */
int gsb_compute_stream(gsb_t* h, void** stream);

/* ---- stateless kernel entry for callers that own device arrays ------------------------- */

/* One calculate! on DEVICE arrays in the reference layout ((sx+2)(sy+2)(sz+2), unpadded), e.g.
 * CuArrays owned by the caller (the shape ext/CUDAExt.jl:175-186 launches on).  Writes the
 * interior of u_out/v_out.  offsets = global coordinates of the first interior cell (noise
 * key); step = noise counter.  stream: a cudaStream_t or NULL. */
int gsb_calculate_arrays(int dtype, int arith, const int64_t sizes[3], const int64_t offsets[3],
                         const void* u_dev, const void* v_dev, void* u_out_dev, void* v_out_dev,
                         double Du, double Dv, double F, double k, double dt, double noise,
                         uint64_t seed, uint32_t step, void* stream);

/* ---- statistics without any allocation --------------------------------------------------- */

/* No counterpart in the reference (it has no on-device reductions; its analysis script works on
 * the files ADIOS2 wrote).
 *
 * Same five numbers as gsb_field_stats -- sum, sum of squares, min, max, count(value > thresh)
 * over the interior cells -- but for a caller that may not allocate device memory: a host whose
 * GPU memory is owned end to end by another allocator (CUDA.jl's pool, PyTorch's caching
 * allocator), where a cudaMalloc from inside this library can fail or fragment although the pool
 * has room, or an in-situ hook inside the time loop, where cudaMalloc/cudaFree could synchronise
 * the device and stall the overlapped halo exchange.  No handle is needed (gsb_create allocates).
 *
 *   field_dev : DEVICE array in the reference layout, ghosted: (sx+2)(sy+2)(sz+2), x fastest
 *   out_dev   : 5 doubles the GPU can write (device or pinned host memory), caller-owned
 *   stream    : a cudaStream_t or NULL; the call only enqueues two kernels
 *   errbuf    : optional; receives the reason when the result is non-zero (gsb_last_error()
 *               returns the same message)
 * Like step/exchange/calculate it only enqueues: out_dev is valid once `stream` has completed.
 *
 * The per-plane partial results go to a fixed array inside the library image (160 KiB of device
 * memory on every GPU that loads the library, held until exit), so:
 *   - sizes[2] (z-planes) must be <= GSB_NOALLOC_MAX_PLANES, larger fields are refused;
 *   - calls that overlap on the GPU (different streams, no ordering) race on that array -- keep
 *     them on one stream or order them with events; back-to-back calls on one stream are fine
 *     and need no clearing in between (every slot that is read is written first, each call);
 *   - the FIRST call on each device resolves the array's device address and should therefore be
 *     made before any stream capture begins on that device.
 * It calls no cudaMalloc / cudaFree / allocating library routine. */
#define GSB_NOALLOC_MAX_PLANES 4096
int gsb_field_stats_noalloc(int dtype, const int64_t sizes[3], const void* field_dev, double thresh,
                            double* out_dev, void* stream, char* errbuf, size_t errbuf_len);

#ifdef __cplusplus
}
#endif
#endif /* GSB200_H */
