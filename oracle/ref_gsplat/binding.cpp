// TEST INFRASTRUCTURE (not product code): exposes the REFERENCE's own CUDA rasterization kernels
// (vendored gsplat 1.5.3, compiled in place from /root/reference by oracle/build_ref.sh for sm_100a)
// to the GPU parity tests and to scripts/compare_gsplat.py, so the B200-native blend kernels can be
// checked against -- and timed next to -- the kernels the reference actually runs
// (cpp/third_party/gsplat/gsplat/cuda/csrc/RasterizeToPixels3DGS{Fwd,Bwd}.cu).  No reference source
// is copied: this file only declares the launchers of Rasterization.h and calls them.
#include <torch/extension.h>

#include "Common.h"
#include "Rasterization.h"

namespace {

template <uint32_t C>
std::vector<at::Tensor> fwd_c(const at::Tensor& means2d, const at::Tensor& conics, const at::Tensor& colors,
                              const at::Tensor& opac, int64_t W, int64_t H, const at::Tensor& offsets,
                              const at::Tensor& ids) {
    auto o = means2d.options();
    at::Tensor renders = at::empty({1, H, W, (int64_t)C}, o);
    at::Tensor alphas = at::empty({1, H, W, 1}, o);
    at::Tensor last = at::empty({1, H, W}, o.dtype(at::kInt));
    gsplat::launch_rasterize_to_pixels_3dgs_fwd_kernel<C>(means2d, conics, colors, opac, c10::nullopt, c10::nullopt,
                                                          (uint32_t)W, (uint32_t)H, 16, offsets, ids, renders, alphas,
                                                          last);
    return {renders, alphas, last};
}

template <uint32_t C>
std::vector<at::Tensor> bwd_c(const at::Tensor& means2d, const at::Tensor& conics, const at::Tensor& colors,
                              const at::Tensor& opac, int64_t W, int64_t H, const at::Tensor& offsets,
                              const at::Tensor& ids, const at::Tensor& alphas, const at::Tensor& last,
                              const at::Tensor& v_colors, const at::Tensor& v_alphas) {
    at::Tensor v_means2d = at::zeros_like(means2d);
    at::Tensor v_conics = at::zeros_like(conics);
    at::Tensor v_cols = at::zeros_like(colors);
    at::Tensor v_opac = at::zeros_like(opac);
    gsplat::launch_rasterize_to_pixels_3dgs_bwd_kernel<C>(means2d, conics, colors, opac, c10::nullopt, c10::nullopt,
                                                          (uint32_t)W, (uint32_t)H, 16, offsets, ids, alphas, last,
                                                          v_colors, v_alphas, c10::nullopt, v_means2d, v_conics, v_cols,
                                                          v_opac);
    return {v_means2d, v_conics, v_cols, v_opac};
}

std::vector<at::Tensor> rast_fwd(const at::Tensor& means2d, const at::Tensor& conics, const at::Tensor& colors,
                                 const at::Tensor& opac, int64_t W, int64_t H, const at::Tensor& offsets,
                                 const at::Tensor& ids) {
    const int64_t C = colors.size(-1);
    if (C == 4) return fwd_c<4>(means2d, conics, colors, opac, W, H, offsets, ids);
    if (C == 5) return fwd_c<5>(means2d, conics, colors, opac, W, H, offsets, ids);
    if (C == 3) return fwd_c<3>(means2d, conics, colors, opac, W, H, offsets, ids);
    TORCH_CHECK(false, "channels must be 3, 4 or 5");
}

std::vector<at::Tensor> rast_bwd(const at::Tensor& means2d, const at::Tensor& conics, const at::Tensor& colors,
                                 const at::Tensor& opac, int64_t W, int64_t H, const at::Tensor& offsets,
                                 const at::Tensor& ids, const at::Tensor& alphas, const at::Tensor& last,
                                 const at::Tensor& v_colors, const at::Tensor& v_alphas) {
    const int64_t C = colors.size(-1);
    if (C == 4) return bwd_c<4>(means2d, conics, colors, opac, W, H, offsets, ids, alphas, last, v_colors, v_alphas);
    if (C == 5) return bwd_c<5>(means2d, conics, colors, opac, W, H, offsets, ids, alphas, last, v_colors, v_alphas);
    if (C == 3) return bwd_c<3>(means2d, conics, colors, opac, W, H, offsets, ids, alphas, last, v_colors, v_alphas);
    TORCH_CHECK(false, "channels must be 3, 4 or 5");
}

} // namespace

PYBIND11_MODULE(gsplat_ref, m) {
    m.doc() = "reference gsplat rasterization kernels (test infrastructure)";
    m.def("rast_fwd", &rast_fwd);
    m.def("rast_bwd", &rast_bwd);
}
