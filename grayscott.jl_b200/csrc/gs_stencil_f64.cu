// Float64 instantiations of the stencil kernels (see gs_stencil.cuh).
#include "gs_stencil_launch.inc"
namespace gsb {
template cudaError_t launch_stencil_T<double>(const StencilCfg&, const void*, const void*, void*, void*,
                                              const CUtensorMap*, const CUtensorMap*, const Layout&, const Params&,
                                              const Box&, const XFaceArgs&, cudaStream_t);
template cudaError_t launch_naive_T<double>(int, int, const void*, const void*, void*, void*, int, int, int,
                                            const Params&, cudaStream_t);
}  // namespace gsb
