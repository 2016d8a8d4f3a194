"""`CubemapEncoder` / `MipCubemapEncoder` over the sm_100a lookup kernels.

Same constructor arguments, attributes, parameter names (`params.Cubemap_texture`,
`params.Cubemap_failv` -- the keys scene/gaussian_model.py:334 saves and loads), methods
(`resize`, `filter`, `set_textures`) and output layout ([B,C]) as the reference's
C/cubemapencoder/cubemap_encoder.py:19-178."""
import math

import torch
import torch.nn as nn

import _cubemapencoder as _backend

_interp_to_id = {"nearest": 0, "linear": 1}


class _CubemapEncode(torch.autograd.Function):
    """reference: _cubemap_encode, cubemap_encoder.py:19-63 (forward casts to fp32 like custom_fwd)."""

    @staticmethod
    def forward(ctx, inputs, embeddings, fail_value, interpolation, enable_seamless):
        if torch.is_autocast_enabled():
            inputs, embeddings, fail_value = inputs.float(), embeddings.float(), fail_value.float()
        with torch.autocast("cuda", enabled=False):
            embeddings = embeddings.contiguous()
            inputs = inputs.contiguous()
            fail_value = fail_value.contiguous()
            Cn, L, B = embeddings.shape[1], embeddings.shape[2], inputs.shape[0]
            outputs = torch.empty([Cn, B], dtype=embeddings.dtype, device=embeddings.device)
            _backend.cubemap_encode_forward(inputs, embeddings, fail_value, outputs, interpolation, enable_seamless, B, Cn, L)
        ctx.save_for_backward(inputs, embeddings)
        ctx.mode = (int(interpolation), int(enable_seamless))
        return outputs

    @staticmethod
    def backward(ctx, grad_outputs):
        inputs, embeddings = ctx.saved_tensors
        grad_outputs = grad_outputs.contiguous()
        Cn, L, B = embeddings.shape[1], embeddings.shape[2], inputs.shape[0]
        grad_embeddings = torch.zeros_like(embeddings)
        grad_inputs = torch.empty_like(inputs)
        grad_fail = torch.zeros([Cn], dtype=embeddings.dtype, device=embeddings.device)
        _backend.cubemap_encode_backward(grad_outputs, inputs, embeddings, grad_embeddings, grad_inputs, grad_fail,
                                         ctx.mode[0], ctx.mode[1], B, Cn, L)
        return grad_inputs, grad_embeddings, grad_fail, None, None


cubemap_encode = _CubemapEncode.apply


class CubemapEncoder(nn.Module):
    def __init__(self, output_dim=6, resolution=256, interpolation="linear"):
        super().__init__()
        self.input_dim = 3
        self.resolution = resolution
        self.output_dim = output_dim
        self.interpolation = interpolation
        self.interp_id = _interp_to_id[interpolation]
        self.seamless = 1
        self.params = nn.ParameterDict({
            "Cubemap_texture": nn.Parameter(torch.rand(6, self.output_dim, resolution, resolution) - 0.5),
            "Cubemap_failv": nn.Parameter(torch.zeros(self.output_dim)),
        })
        self.n_elems = 6 * self.output_dim * resolution * resolution + self.output_dim

    def __repr__(self):
        return (f"CubemapEncoder: input_dim={self.input_dim} output_dim={self.output_dim} resolution={self.resolution} "
                f"-> {self.n_elems} interpolation={self.interpolation} seamless={self.seamless}")

    def resize(self, new_resolution):
        self.resolution = new_resolution
        self.params["Cubemap_texture"] = nn.functional.interpolate(
            self.params["Cubemap_texture"], size=(new_resolution, new_resolution), mode="bicubic", align_corners=True)
        self.n_elems = 6 * self.output_dim * self.resolution * self.resolution + self.output_dim

    def filter(self, activation, inverse_activation, factor=2.0):
        from torchvision.transforms import functional as F   # only this method needs torchvision
        textures = activation(self.params["Cubemap_texture"])
        textures = F.adjust_sharpness(textures, factor)
        textures = torch.clamp(textures, min=1e-3, max=1 - 1e-3)
        self.params["Cubemap_texture"] = inverse_activation(textures)

    def set_textures(self, textures):
        self.resolution = textures.shape[2]
        self.params["Cubemap_texture"] = nn.Parameter(textures)
        self.n_elems = 6 * self.output_dim * self.resolution * self.resolution + self.output_dim

    def forward(self, inputs):
        outputs = cubemap_encode(inputs, self.params["Cubemap_texture"], self.params["Cubemap_failv"], self.interp_id,
                                 self.seamless)
        return outputs.permute(1, 0)   # [C,B] -> [B,C]


class MipCubemapEncoder(nn.Module):
    def __init__(self, num_levels=4, level_dim=6, per_level_scale=4, base_resolution=4, interpolation="linear",
                 concat=True):
        super().__init__()
        self.input_dim = 3
        self.num_levels = num_levels
        self.level_dim = level_dim
        self.per_level_scale = per_level_scale
        self.base_resolution = base_resolution
        self.concat = concat
        self.output_dim = num_levels * level_dim if concat else level_dim
        self.interpolation = interpolation
        self.interp_id = _interp_to_id[interpolation]
        self.seamless = 1
        levels, n_elems, L = [], 0, float(base_resolution)
        for _ in range(num_levels):
            iL = int(math.ceil(L))
            levels.append(nn.Parameter(torch.empty(6, self.level_dim, iL, iL)))
            n_elems += 6 * self.level_dim * iL * iL
            L *= per_level_scale
        self.params_list = nn.ParameterList(levels)
        self.fail_value = nn.Parameter(torch.zeros(self.level_dim))
        self.n_elems = n_elems + self.level_dim
        self.reset_parameters()

    def reset_parameters(self):
        for p in self.params_list:
            p.data.uniform_(-1e-4, 1e-4)

    def __repr__(self):
        return (f"MipCubemapEncoder: input_dim={self.input_dim} num_levels={self.num_levels} level_dim={self.level_dim} "
                f"base_resolution={self.base_resolution} -> {self.n_elems} per_level_scale={self.per_level_scale:.4f} "
                f"interpolation={self.interpolation} seamless={self.seamless}")

    def forward(self, inputs):
        outs = [cubemap_encode(inputs, p, self.fail_value, self.interp_id, self.seamless) for p in self.params_list]
        outputs = torch.cat(outs, dim=0) if self.concat else sum(outs)
        return outputs.permute(1, 0)
