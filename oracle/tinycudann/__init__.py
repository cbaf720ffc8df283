"""ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product path (loner_b200/).

Pure-torch fp32 restatement of the three `tinycudann` modules the reference instantiates
(reference call sites: src/models/nerf_tcnn.py:35-38 `NetworkWithInputEncoding`, :40,:43
`Encoding`, :50-52 `Network`).  The arithmetic lives in the third-party dependency
NVlabs/tiny-cuda-nn (PyTorch bindings `tinycudann`, version 1.7 per reference README.md:343,
installed from un-pinned git HEAD in docker/container_dockerhub.Dockerfile:64).  Its source is NOT
under /root/reference and it cannot be installed here, so this file restates the *published*
algorithm (Mueller et al., "Instant Neural Graphics Primitives", and the public tiny-cuda-nn
`encodings/grid.h`, `encodings/spherical_harmonics.h`, `networks/fully_fused_mlp.cu` semantics):

  * level scale  s_l = 2^(l*log2(per_level_scale)) * base_resolution - 1,  res_l = ceil(s_l) + 1
  * params/level = min(roundup8(res^3 capped at 2^31-1), 2^log2_hashmap_size), offsets cumulative
  * position     p = s_l * x + 0.5,  cell = floor(p),  frac = p - cell,  trilinear weights
  * index        dense x + y*res + z*res^2 while the running stride never exceeds the level size,
                 else (x*1) ^ (y*2654435761) ^ (z*805459861) in uint32; finally `% size_l`
  * features     y[l*F + f] = sum_c w_c * table[(off_l + idx_c)*F + f]
  * MLP          no biases, ReLU hidden, row-major [out, in] matrices, output padded to 16
  * param vector network weights first, then encoding tables (NetworkWithInputEncoding)
  * init         tables U(-1e-4, 1e-4); MLP Xavier-uniform

PARITY STATUS: **unpinned at the tcnn boundary** -- the reference has no test or golden vector for
these modules and no tcnn binary exists in this environment.  Everything *around* them (sampling,
compositing, loss) is pinned against the reference's own Python (see oracle/make_golden.py).
This shim computes in fp32 throughout (tcnn computes in fp16); north_star asks for fp32 accumulate.
"""
import math

import numpy as np
import torch
import torch.nn as nn

PRIMES = (1, 2654435761, 805459861)


def level_table(n_levels, base_resolution, log2_hashmap_size, per_level_scale=2.0, n_dims=3):
    """Returns list of dicts {scale, res, offset, size, dense} following tcnn GridEncoding ctor."""
    out = []
    offset = 0
    log2_pls = math.log2(per_level_scale)
    for l in range(n_levels):
        scale = float(np.float32(np.exp2(np.float32(l * log2_pls))) * np.float32(base_resolution) - np.float32(1.0))
        res = int(math.ceil(scale)) + 1
        max_params = (2 ** 32 - 1) // 2
        dense_params = res ** n_dims
        params = max_params if float(res) ** n_dims > float(max_params) else dense_params
        params = (params + 7) // 8 * 8
        params = min(params, 1 << log2_hashmap_size)
        # dense iff the running stride never exceeds the level size (grid_index loop)
        stride, is_dense = 1, True
        for _ in range(n_dims):
            if stride > params:
                break
            stride *= res
        is_dense = not (params < stride)
        out.append(dict(scale=scale, res=res, offset=offset, size=params, dense=is_dense))
        offset += params
    return out, offset


def grid_indices(ix, iy, iz, lvl):
    """ix,iy,iz: int64 tensors of corner coords. Returns int64 index into the level's table."""
    size, res = lvl["size"], lvl["res"]
    M = 0xFFFFFFFF
    if lvl["dense"]:
        idx = (ix + iy * res + iz * res * res) & M
    else:
        idx = ((ix * PRIMES[0]) & M) ^ ((iy * PRIMES[1]) & M) ^ ((iz * PRIMES[2]) & M)
    return idx % size


class _HashGridFn:
    """Differentiable (torch autograd, fp32) multi-resolution hash encoding."""

    def __init__(self, cfg):
        self.n_levels = int(cfg["n_levels"])
        self.F = int(cfg["n_features_per_level"])
        self.levels, self.n_entries = level_table(
            self.n_levels, int(cfg["base_resolution"]), int(cfg["log2_hashmap_size"]),
            float(cfg.get("per_level_scale", 2.0)))
        self.n_params = self.n_entries * self.F
        self.n_output_dims = self.n_levels * self.F

    def __call__(self, x, table_flat):
        # x [P,3] in [0,1]; table_flat [n_entries*F]
        table = table_flat.view(self.n_entries, self.F)
        outs = []
        for lvl in self.levels:
            # tcnn pos_fract: fmaf(scale, x, 0.5f) -- ONE rounding.  At the finest levels p ~ 5e5 has an
            # ulp of 2^-5, so a doubly-rounded x*scale+0.5 would move the fractional position by up to
            # 3 % of a cell; fp64 evaluation rounded once to fp32 reproduces the fused op exactly.
            p = (x.double() * lvl["scale"] + 0.5).float()
            cell = torch.floor(p)
            frac = p - cell
            ci = cell.detach().to(torch.int64)
            acc = 0
            for c in range(8):
                bx, by, bz = c & 1, (c >> 1) & 1, (c >> 2) & 1
                wx = frac[:, 0] if bx else 1 - frac[:, 0]
                wy = frac[:, 1] if by else 1 - frac[:, 1]
                wz = frac[:, 2] if bz else 1 - frac[:, 2]
                w = wx * wy * wz
                idx = grid_indices(ci[:, 0] + bx, ci[:, 1] + by, ci[:, 2] + bz, lvl) + lvl["offset"]
                acc = acc + w.unsqueeze(-1) * table[idx]
            outs.append(acc)
        return torch.cat(outs, dim=-1)


def sh4(d01):
    """Degree-4 real spherical harmonics (16 coeffs) on 2*d01-1 (tcnn SphericalHarmonics)."""
    v = d01 * 2 - 1
    x, y, z = v[:, 0], v[:, 1], v[:, 2]
    xy, xz, yz, x2, y2, z2 = x * y, x * z, y * z, x * x, y * y, z * z
    o = [
        torch.full_like(x, 0.28209479177387814),
        -0.48860251190291987 * y,
        0.48860251190291987 * z,
        -0.48860251190291987 * x,
        1.0925484305920792 * xy,
        -1.0925484305920792 * yz,
        0.94617469575755997 * z2 - 0.31539156525251999,
        -1.0925484305920792 * xz,
        0.54627421529603959 * x2 - 0.54627421529603959 * y2,
        0.59004358992664352 * y * (-3.0 * x2 + y2),
        2.8906114426405538 * xy * z,
        0.45704579946446572 * y * (1.0 - 5.0 * z2),
        0.3731763325901154 * z * (5.0 * z2 - 3.0),
        0.45704579946446572 * x * (1.0 - 5.0 * z2),
        1.4453057213202769 * z * (x2 - y2),
        0.59004358992664352 * x * (-x2 + 3.0 * y2),
    ]
    return torch.stack(o, dim=-1)


def _pad16(n):
    return (n + 15) // 16 * 16


class _MLP:
    """FullyFusedMLP semantics: widths in -> n_neurons x n_hidden -> pad16(out); no bias."""

    def __init__(self, n_in, n_out, cfg):
        self.n_in = _pad16(n_in)
        self.n_in_raw = n_in
        self.n_out = n_out
        self.n_out_padded = _pad16(n_out)
        self.width = int(cfg["n_neurons"])
        self.n_hidden = int(cfg["n_hidden_layers"])
        act = cfg.get("activation", "ReLU")
        oact = cfg.get("output_activation", "None")
        assert act == "ReLU", "oracle restates ReLU only (reference cfg/nerf_config/default_nerf_hash.yaml)"
        assert str(oact) == "None"
        self.shapes = [(self.width, self.n_in)]
        for _ in range(self.n_hidden - 1):
            self.shapes.append((self.width, self.width))
        self.shapes.append((self.n_out_padded, self.width))
        self.n_params = sum(a * b for a, b in self.shapes)

    def init(self, gen):
        chunks = []
        for (o, i) in self.shapes:
            s = math.sqrt(6.0 / (i + o))
            chunks.append((torch.rand(o * i, generator=gen) * 2 - 1) * s)
        return torch.cat(chunks)

    def __call__(self, x, params):
        if x.shape[1] < self.n_in:
            x = torch.nn.functional.pad(x, (0, self.n_in - x.shape[1]), value=1.0)
        off = 0
        h = x
        for li, (o, i) in enumerate(self.shapes):
            W = params[off:off + o * i].view(o, i)
            off += o * i
            h = h @ W.t()
            if li < len(self.shapes) - 1:
                h = torch.relu(h)
        return h[:, :self.n_out]


class Encoding(nn.Module):
    def __init__(self, n_input_dims, encoding_config, seed=1337, dtype=None):
        super().__init__()
        assert n_input_dims == 3
        self.otype = encoding_config["otype"]
        self.dtype = torch.float32
        if self.otype == "HashGrid":
            self._fn = _HashGridFn(encoding_config)
            self.n_output_dims = self._fn.n_output_dims
            g = torch.Generator().manual_seed(seed)
            init = (torch.rand(self._fn.n_params, generator=g) * 2 - 1) * 1e-4
            self.params = nn.Parameter(init)
        elif self.otype == "SphericalHarmonics":
            assert int(encoding_config["degree"]) == 4
            self.n_output_dims = 16
            self.params = nn.Parameter(torch.zeros(0))
        else:
            raise NotImplementedError(self.otype)

    def forward(self, x):
        if self.otype == "HashGrid":
            return self._fn(x.float(), self.params)
        return sh4(x.float())


class Network(nn.Module):
    def __init__(self, n_input_dims, n_output_dims, network_config, seed=1337):
        super().__init__()
        self.dtype = torch.float32
        self._mlp = _MLP(n_input_dims, n_output_dims, network_config)
        self.n_output_dims = n_output_dims
        g = torch.Generator().manual_seed(seed)
        self.params = nn.Parameter(self._mlp.init(g))

    def forward(self, x):
        return self._mlp(x.float(), self.params)


class NetworkWithInputEncoding(nn.Module):
    def __init__(self, n_input_dims, n_output_dims, encoding_config, network_config, seed=1337):
        super().__init__()
        assert n_input_dims == 3 and encoding_config["otype"] == "HashGrid"
        self.dtype = torch.float32
        self._enc = _HashGridFn(encoding_config)
        self._mlp = _MLP(self._enc.n_output_dims, n_output_dims, network_config)
        self.n_output_dims = n_output_dims
        g = torch.Generator().manual_seed(seed)
        mlp = self._mlp.init(g)
        grid = (torch.rand(self._enc.n_params, generator=g) * 2 - 1) * 1e-4
        # parameter order: network weights, then encoding tables
        self.params = nn.Parameter(torch.cat([mlp, grid]))

    def encode(self, x):
        return self._enc(x.float(), self.params[self._mlp.n_params:])

    def forward(self, x):
        e = self.encode(x)
        return self._mlp(e, self.params[:self._mlp.n_params])
