"""PyTorch operator boundary of the rasterizer.

Mirrors the reference's torch binding so an INRIA-style trainer can swap its import:

    reference                                                     here
    lib/diff-gaussian-rasterization/rasterize_points.cu:35-217    class _C (three functions, same tuples)
    lib/diff-gaussian-rasterization/ext.cpp:15-18                 _C.rasterize_gaussians / _backward / mark_visible
    .../diff_gaussian_rasterization/__init__.py:20-220            rasterize_gaussians, _RasterizeGaussians,
                                                                  GaussianRasterizationSettings, GaussianRasterizer

Every call goes through the C ABI (include/gsb.h: gsb_rasterizer_forward / backward / mark_visible)
into the sm_100a kernels; torch supplies device memory, the current stream and autograd.  As in the
reference, an absent optional input is an EMPTY tensor (`torch.Tensor([])`), which reaches the
rasterizer as a null pointer (rasterize_points.cu:96-113), the three workspaces are uint8 tensors
grown by the allocator callbacks (rasterize_points.cu:27-33), and gradient tensors are zero-filled
before the call (rasterize_points.cu:146-154).  There is no CPU path: CPU tensors raise.
"""
import ctypes as C
from typing import NamedTuple

import torch
import torch.nn as nn

from ._lib import ALLOC_FN, check, lib

NUM_CHANNELS = 3            # cuda_rasterizer/config.h:15


def _null_or_ptr(t, name):
    """Empty tensor -> NULL (the reference passes .data_ptr() of an empty tensor, which is null)."""
    if t is None or t.numel() == 0:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: this rasterizer has no CPU path")
    if t.dtype != torch.float32 and t.dtype != torch.int32 and t.dtype != torch.uint8 and t.dtype != torch.bool:
        raise RuntimeError(f"{name}: unsupported dtype {t.dtype}")
    return C.c_void_p(t.data_ptr())


def _f32(t):
    """`.contiguous().data<float>()` of the reference: contiguous float32 view (no copy when it already is)."""
    if t is None or t.numel() == 0:
        return t
    if t.dtype != torch.float32:
        raise RuntimeError(f"expected a float32 tensor, got {t.dtype}")
    return t.contiguous()


def _resize_functional(holder, key, device):
    """resizeFunctional (rasterize_points.cu:27-33): grow the uint8 tensor, hand back its pointer."""
    def alloc(nbytes, _user):
        t = holder[key]
        t.resize_(int(nbytes))
        return t.data_ptr() if nbytes else 0
    return ALLOC_FN(alloc)


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class _C:
    """The three functions ext.cpp:15-18 exports, with rasterize_points.cu's argument order and tuples."""

    @staticmethod
    def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                            viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                            prefiltered, debug):
        """RasterizeGaussiansCUDA (rasterize_points.cu:35-118) ->
        (num_rendered, out_color[3,H,W], radii[P], geomBuffer, binningBuffer, imgBuffer)."""
        if means3D.dim() != 2 or means3D.size(1) != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")
        P, H, W = means3D.size(0), int(image_height), int(image_width)
        if P != 0 and not means3D.is_cuda:
            raise RuntimeError("means3D must be a CUDA tensor: this rasterizer has no CPU path")
        dev = means3D.device
        out_color = torch.zeros((NUM_CHANNELS, H, W), dtype=torch.float32, device=dev)
        radii = torch.zeros((P,), dtype=torch.int32, device=dev)
        bufs = {k: torch.empty((0,), dtype=torch.uint8, device=dev) for k in ("geom", "binning", "img")}
        rendered = 0
        if P != 0:
            M = sh.size(1) if sh.numel() != 0 and sh.size(0) != 0 else 0
            cbs = [_resize_functional(bufs, k, dev) for k in ("geom", "binning", "img")]
            keep = [_f32(t) for t in (background, means3D, sh, colors, opacity, scales, rotations, cov3D_precomp,
                                      viewmatrix, projmatrix, campos)]
            bg, m3, shc, col, opa, sca, rot, cov, vm, pm, cp = keep
            n = C.c_int(0)
            with torch.cuda.device(dev):
                check(lib().gsb_rasterizer_forward(
                    cbs[0], None, cbs[1], None, cbs[2], None, P, int(degree), M, _null_or_ptr(bg, "background"), W, H,
                    _null_or_ptr(m3, "means3D"), _null_or_ptr(shc, "sh"), _null_or_ptr(col, "colors_precomp"),
                    _null_or_ptr(opa, "opacities"), _null_or_ptr(sca, "scales"), float(scale_modifier),
                    _null_or_ptr(rot, "rotations"), _null_or_ptr(cov, "cov3D_precomp"), _null_or_ptr(vm, "viewmatrix"),
                    _null_or_ptr(pm, "projmatrix"), _null_or_ptr(cp, "campos"), float(tan_fovx), float(tan_fovy),
                    int(bool(prefiltered)), C.c_void_p(out_color.data_ptr()), C.c_void_p(radii.data_ptr()),
                    int(bool(debug)), _stream(dev), C.byref(n)))
            rendered = n.value
        return rendered, out_color, radii, bufs["geom"], bufs["binning"], bufs["img"]

    @staticmethod
    def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                     cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh,
                                     degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug):
        """RasterizeGaussiansBackwardCUDA (rasterize_points.cu:120-194) ->
        (dL_dmeans2D[P,3], dL_dcolors[P,3], dL_dopacity[P,1], dL_dmeans3D[P,3], dL_dcov3D[P,6], dL_dsh[P,M,3],
         dL_dscales[P,3], dL_drotations[P,4])."""
        P = means3D.size(0)
        H, W = dL_dout_color.size(1), dL_dout_color.size(2)
        M = sh.size(1) if sh.numel() != 0 and sh.size(0) != 0 else 0
        o = dict(dtype=torch.float32, device=means3D.device)
        dL_dmeans3D = torch.zeros((P, 3), **o)
        dL_dmeans2D = torch.zeros((P, 3), **o)
        dL_dcolors = torch.zeros((P, NUM_CHANNELS), **o)
        dL_dconic = torch.zeros((P, 2, 2), **o)
        dL_dopacity = torch.zeros((P, 1), **o)
        dL_dcov3D = torch.zeros((P, 6), **o)
        dL_dsh = torch.zeros((P, M, 3), **o)
        dL_dscales = torch.zeros((P, 3), **o)
        dL_drotations = torch.zeros((P, 4), **o)
        if P != 0:
            dev = means3D.device
            keep = [_f32(t) for t in (background, means3D, sh, colors, scales, rotations, cov3D_precomp, viewmatrix,
                                      projmatrix, campos, dL_dout_color)]
            bg, m3, shc, col, sca, rot, cov, vm, pm, cp, dpix = keep
            rad = radii.contiguous()
            p = lambda t: C.c_void_p(t.data_ptr())
            with torch.cuda.device(dev):
                check(lib().gsb_rasterizer_backward(
                    P, int(degree), M, int(R), _null_or_ptr(bg, "background"), W, H, _null_or_ptr(m3, "means3D"),
                    _null_or_ptr(shc, "sh"), _null_or_ptr(col, "colors_precomp"), _null_or_ptr(sca, "scales"),
                    float(scale_modifier), _null_or_ptr(rot, "rotations"), _null_or_ptr(cov, "cov3D_precomp"),
                    _null_or_ptr(vm, "viewmatrix"), _null_or_ptr(pm, "projmatrix"), _null_or_ptr(cp, "campos"),
                    float(tan_fovx), float(tan_fovy), _null_or_ptr(rad, "radii"), _null_or_ptr(geomBuffer, "geomBuffer"),
                    _null_or_ptr(binningBuffer, "binningBuffer"), _null_or_ptr(imageBuffer, "imageBuffer"),
                    _null_or_ptr(dpix, "dL_dout_color"), p(dL_dmeans2D), p(dL_dconic), p(dL_dopacity), p(dL_dcolors),
                    p(dL_dmeans3D), p(dL_dcov3D), p(dL_dsh) if M else None, p(dL_dscales), p(dL_drotations),
                    int(bool(debug)), _stream(dev)))
        return dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations

    @staticmethod
    def mark_visible(means3D, viewmatrix, projmatrix):
        """markVisible (rasterize_points.cu:196-217) -> bool[P]."""
        P = means3D.size(0)
        present = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
        if P != 0:
            if not means3D.is_cuda:
                raise RuntimeError("means3D must be a CUDA tensor: this rasterizer has no CPU path")
            m3, vm, pm = _f32(means3D), _f32(viewmatrix), _f32(projmatrix)
            with torch.cuda.device(means3D.device):
                check(lib().gsb_rasterizer_mark_visible(P, C.c_void_p(m3.data_ptr()), C.c_void_p(vm.data_ptr()),
                                                        C.c_void_p(pm.data_ptr()), C.c_void_p(present.data_ptr()),
                                                        _stream(means3D.device)))
        return present


def cpu_deep_copy_tuple(input_tuple):
    """__init__.py:17-19."""
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    """__init__.py:21-42."""
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class _RasterizeGaussians(torch.autograd.Function):
    """__init__.py:44-152: same argument restructuring, saved tensors and gradient tuple order."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        rs = raster_settings
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh,
                rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args)      # copy them before they can be corrupted
            try:
                num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _C.rasterize_gaussians(*args)
            except Exception:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise
        else:
            num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _C.rasterize_gaussians(*args)
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.opacity_shape = tuple(opacities.shape)     # the reference returns [P,1]; a [P] input gets [P] here
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
                              binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _):
        num_rendered = ctx.num_rendered
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer, binningBuffer,
         imgBuffer) = ctx.saved_tensors
        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color, sh, rs.sh_degree, rs.campos,
                geomBuffer, num_rendered, binningBuffer, imgBuffer, rs.debug)
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                out = _C.rasterize_gaussians_backward(*args)
            except Exception:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise
        else:
            out = _C.rasterize_gaussians_backward(*args)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations) = out
        # autograd wants None (or a matching shape) for inputs that were passed as empty tensors
        fit = lambda g, x: g if x.numel() == g.numel() and x.numel() != 0 else None
        return (grad_means3D, grad_means2D, fit(grad_sh, sh), fit(grad_colors_precomp, colors_precomp),
                grad_opacities.reshape(ctx.opacity_shape), fit(grad_scales, scales),
                fit(grad_rotations, rotations), fit(grad_cov3Ds_precomp, cov3Ds_precomp), None)


class GaussianRasterizationSettings(NamedTuple):
    """__init__.py:154-166."""
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class GaussianRasterizer(nn.Module):
    """__init__.py:168-220."""

    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        rs = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   rs)
