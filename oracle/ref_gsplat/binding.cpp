// TEST INFRASTRUCTURE (not product code): exposes the REFERENCE's own CUDA rasterization kernels
// (vendored gsplat 1.5.3, compiled in place from /root/reference by oracle/build_ref.sh for sm_100a)
// to the GPU parity tests and to scripts/compare_gsplat.py, so the B200-native blend kernels can be
// checked against -- and timed next to -- the kernels the reference actually runs
// (cpp/third_party/gsplat/gsplat/cuda/csrc/RasterizeToPixels3DGS{Fwd,Bwd}.cu).  No reference source
// is copied: this file only declares the launchers of Rasterization.h and calls them.
#include <torch/extension.h>

#include "Common.h"
#include "Intersect.h"
#include "Ops.h"
#include "Projection.h"
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

// projection_ewa_3dgs_fused_fwd (Projection.cpp:94-160 allocates the outputs and calls this launcher; that file also
// pulls in every other projection variant, so only the launcher of ProjectionEWA3DGSFused.cu is linked).
// means [N,3], quats [N,4], scales [N,3], opacities [N] or undefined (=> fixed 3.33 sigma extent), viewmat [4,4], K [3,3].
std::vector<at::Tensor> proj_fwd(const at::Tensor& means, const at::Tensor& quats, const at::Tensor& scales,
                                 const at::Tensor& opacities, const at::Tensor& viewmat, const at::Tensor& K, int64_t W,
                                 int64_t H, double eps2d, double near_plane, double far_plane) {
    const int64_t N = means.size(0);
    auto o = means.options();
    at::Tensor radii = at::empty({1, N, 2}, o.dtype(at::kInt));
    at::Tensor means2d = at::empty({1, N, 2}, o);
    at::Tensor depths = at::empty({1, N}, o);
    at::Tensor conics = at::empty({1, N, 3}, o);
    at::optional<at::Tensor> opac = opacities.defined() ? at::optional<at::Tensor>(opacities) : c10::nullopt;
    gsplat::launch_projection_ewa_3dgs_fused_fwd_kernel(means, c10::nullopt, quats, scales, opac, viewmat.reshape({1, 4, 4}),
                                                        K.reshape({1, 3, 3}), (uint32_t)W, (uint32_t)H, (float)eps2d,
                                                        (float)near_plane, (float)far_plane, 0.f, gsplat::PINHOLE, radii,
                                                        means2d, depths, conics, c10::nullopt);
    return {radii.reshape({N, 2}), means2d.reshape({N, 2}), depths.reshape({N}), conics.reshape({N, 3})};
}

// the reference's own host path: intersect_tile (two passes + cumsum + .item() + CUB radix sort) and intersect_offset
// (Intersect.cpp:15-186), on [1,N,*] inputs, I = 1, tile 16, sorted, unsegmented -- what renderer.cpp:192-212 calls.
std::vector<at::Tensor> isect(const at::Tensor& means2d, const at::Tensor& radii, const at::Tensor& depths, int64_t tile_w,
                              int64_t tile_h) {
    const int64_t N = means2d.size(0);
    auto r = gsplat::intersect_tile(means2d.reshape({1, N, 2}), radii.reshape({1, N, 2}), depths.reshape({1, N}),
                                    c10::nullopt, c10::nullopt, 1, 16, (uint32_t)tile_w, (uint32_t)tile_h, true, false);
    at::Tensor offsets = gsplat::intersect_offset(std::get<1>(r), 1, (uint32_t)tile_w, (uint32_t)tile_h);
    return {std::get<0>(r).reshape({N}), std::get<1>(r), std::get<2>(r), offsets.reshape({tile_h, tile_w})};
}

} // namespace

PYBIND11_MODULE(gsplat_ref, m) {
    m.doc() = "reference gsplat rasterization / projection / intersection kernels (test infrastructure)";
    m.def("rast_fwd", &rast_fwd);
    m.def("rast_bwd", &rast_bwd);
    m.def("proj_fwd", &proj_fwd);
    m.def("isect", &isect);
}
