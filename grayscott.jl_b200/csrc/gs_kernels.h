// gs_kernels.h -- host-side launch interface of the device code (internal; the public
// boundary is include/gsb200.h).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include "gs_common.cuh"

namespace gsb {

enum : int { DT_F32 = 0, DT_F64 = 1 };
enum : int { KERNEL_TMA = 0, KERNEL_L1 = 1, KERNEL_TMA_LO = 2, KERNEL_TMA_HI = 3 };

struct StencilCfg {
    int dtype;     // DT_F32 / DT_F64
    int arith;     // ARITH_REFERENCE / ARITH_NATIVE (F32 only)
    int noise;     // 0/1
    int kind;      // KERNEL_TMA / KERNEL_L1
    int ty;        // rows per CTA
    int stages;    // TMA ring depth
    int zchunk;    // planes per CTA
};

// Valid (ty, stages) for the TMA kernel and ty for the L1 kernel are those instantiated in
// gs_kernels.cu; launch_stencil returns cudaErrorInvalidConfiguration otherwise.
cudaError_t launch_stencil(const StencilCfg& c, const void* u, const void* v, void* uo, void* vo,
                           const CUtensorMap* map_u, const CUtensorMap* map_v, const Layout& L, const Params& p,
                           const Box& box, cudaStream_t st, const XFaceArgs* xface = nullptr);

// Tensor map over one padded field for the TMA kernel with `ty` rows per CTA.  Returns 0 on success.
int make_field_tensor_map(CUtensorMap* out, int dtype, const void* base, const Layout& L, int ty, const char** err);

cudaError_t launch_naive_reflayout(int dtype, int arith, int noise, const void* u, const void* v, void* uo,
                                   void* vo, int sx, int sy, int sz, const Params& p, cudaStream_t st);

// init_fields (Simulation_CPU.jl:14-65): u = 1, v = 0, u_temp = v_temp = 0 everywhere (ghosts and
// padding included), then the 13^3 block around the global centre: u = 0.25, v = 0.33.
// g[d] = global extent, off[d] = this subdomain's 0-based offset.
cudaError_t launch_init_fields(int dtype, void* u, void* v, void* ut, void* vt, const Layout& L,
                               const long long g[3], const long long off[3], cudaStream_t st);

// Halo faces.  dir: 0 west(x-) 1 east(x+) 2 down(y-) 3 up(y+) 4 south(z-) 5 north(z+)
// (proc_neighbors order, communication.jl:90-92).  A face message is [u face | v face], each
// face_elems(dir) elements: x faces sy*sz (y fastest), y faces sx*sz, z faces sx*sy (x fastest) --
// the part of the reference's face datatypes (communication.jl:109-126) a 7-point stencil reads.
struct FaceSet {
    int n;            // number of faces in this launch (<= 6)
    int dir[6];
    void* buf[6];     // message buffer per face (piece)
    int z0[6], z1[6]; // x and y faces: the z range [z0, z1) this piece covers (z faces: ignored)
};
long long face_elems(const Layout& L, int dir);
// elements of the piece of face `dir` restricted to planes [z0, z1) (z faces: the whole plane)
long long face_piece_elems(const Layout& L, int dir, int z0, int z1);
// pack: boundary interior cells of (u, v) -> buf ; unpack: buf -> ghost cells of (u, v)
cudaError_t launch_pack_faces(int dtype, const void* u, const void* v, const Layout& L, const FaceSet& fs,
                              cudaStream_t st);
cudaError_t launch_unpack_faces(int dtype, void* u, void* v, const Layout& L, const FaceSet& fs, cudaStream_t st);

// Per-z-slice histogram of the interior cells (pdfcalc.jl:13-49): pdf_dev[z * nbins + bin], nbins <= 8192.
cudaError_t launch_pdf(int dtype, const void* f, const Layout& L, int nbins, double vmin, double vmax,
                       unsigned* pdf_dev, cudaStream_t st);

// Field statistics over interior cells: out[0]=sum out[1]=sum of squares out[2]=min out[3]=max
// out[4]=count(value > thresh).  `partials` is device scratch of stats_scratch_doubles() doubles;
// `out_dev` 5 doubles on the device.
int stats_scratch_doubles();
cudaError_t launch_field_stats(int dtype, const void* f, const Layout& L, double thresh, double* partials,
                               double* out_dev, cudaStream_t st);

}  // namespace gsb
