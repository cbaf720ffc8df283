"""`tinycudann`-compatible modules backed by libloner_b200.so (NOT a wrapper or port of tiny-cuda-nn).

The reference's DecoupledNeRF (src/models/nerf_tcnn.py:35-52) builds three kinds of tcnn modules:
    tcnn.NetworkWithInputEncoding(3, 1, HashGrid cfg, FullyFusedMLP cfg)     -> sigma head
    tcnn.Encoding(3, HashGrid cfg) / tcnn.Encoding(3, SphericalHarmonics cfg)  -> intensity features
    tcnn.Network(n_in, n_out, FullyFusedMLP cfg)                             -> intensity MLP
This module offers the same constructors, the same single flat fp32 `params` nn.Parameter (network
weights first, then tables -- so reference checkpoints `nerf_model._model_sigma.params` load), the
`.dtype` / `.n_output_dims` attributes the reference reads (nerf_tcnn.py:45-47,54-55), and
`__call__(x[P,3] in [0,1])`.  Compute is fp32 (north_star: fp32 accumulate), so `.dtype` is float32.

`sys.modules['tinycudann'] = loner_b200.tcnn` makes the unmodified reference run on these kernels
(see INTEGRATION.md).
"""
import math

import torch
import torch.nn as nn

from . import ops
from ._lib import LonerB200Error
from .functional import FusedNetwork, HashEncode, SH4, SigmaFromPoints


def _init_tables(n, gen):
    """tcnn grid init: U(-1e-4, 1e-4)."""
    return (torch.rand(n, generator=gen) * 2 - 1) * 1e-4


def _init_mlp(shapes, gen):
    """tcnn FullyFusedMLP init: Xavier-uniform per [out, in] matrix."""
    chunks = []
    for (o, i) in shapes:
        s = math.sqrt(6.0 / (i + o))
        chunks.append((torch.rand(o * i, generator=gen) * 2 - 1) * s)
    return torch.cat(chunks)


def _pad16(n):
    return (n + 15) // 16 * 16


class NetworkWithInputEncoding(nn.Module):
    """Hash grid + fully fused 1-hidden-layer MLP (the sigma head)."""

    def __init__(self, n_input_dims, n_output_dims, encoding_config, network_config, seed=1337):
        super().__init__()
        if n_input_dims != 3 or n_output_dims != 1:
            raise LonerB200Error("NetworkWithInputEncoding: only the 3 -> 1 sigma head is built")
        if network_config.get("otype") != "FullyFusedMLP" or network_config.get("activation") != "ReLU" \
                or str(network_config.get("output_activation")) != "None":
            raise LonerB200Error(f"unsupported network config {dict(network_config)}")
        self.plan = ops.Plan.from_tcnn_config(encoding_config, network_config)
        self.n_input_dims, self.n_output_dims = n_input_dims, n_output_dims
        self.dtype = torch.float32
        self.seed = seed
        g = torch.Generator().manual_seed(seed)
        mlp = _init_mlp([(64, self.plan.n_enc), (16, 64)], g)
        self.params = nn.Parameter(torch.cat([mlp, _init_tables(self.plan.n_table_params, g)]))

    def forward(self, x):
        return SigmaFromPoints.apply(self.plan, self.params, x.float()).unsqueeze(-1)


class Encoding(nn.Module):
    """HashGrid or SphericalHarmonics encoding."""

    def __init__(self, n_input_dims, encoding_config, seed=1337, dtype=None):
        super().__init__()
        if n_input_dims != 3:
            raise LonerB200Error("Encoding: n_input_dims must be 3")
        self.otype = encoding_config.get("otype")
        self.n_input_dims = n_input_dims
        self.dtype = torch.float32
        self.seed = seed
        if self.otype == "HashGrid":
            self.plan = ops.Plan.from_tcnn_config(encoding_config)
            self.n_output_dims = self.plan.n_enc
            g = torch.Generator().manual_seed(seed)
            self.params = nn.Parameter(_init_tables(self.plan.n_table_params, g))
        elif self.otype == "SphericalHarmonics":
            if int(encoding_config.get("degree", 4)) != 4:
                raise LonerB200Error("SphericalHarmonics: only degree 4 (reference default_nerf_hash.yaml:6)")
            self.plan = None
            self.n_output_dims = 16
            self.params = nn.Parameter(torch.zeros(0))
        else:
            raise LonerB200Error(f"unsupported encoding otype {self.otype}")

    def forward(self, x):
        if self.otype == "HashGrid":
            return HashEncode.apply(self.plan, self.params, x.float())
        return SH4.apply(x.float())


class Network(nn.Module):
    """Fully fused MLP without input encoding (the intensity head of the RGB branch, nerf_tcnn.py:50-52):
    n_input_dims (padded with ones to a multiple of 16, as tcnn does) -> 64 x n_hidden_layers -> n_output_dims."""

    def __init__(self, n_input_dims, n_output_dims, network_config, seed=1337):
        super().__init__()
        if network_config.get("otype") != "FullyFusedMLP" or network_config.get("activation") != "ReLU" \
                or int(network_config["n_neurons"]) != 64:
            raise LonerB200Error(f"unsupported network config {dict(network_config)}")
        out_act = str(network_config.get("output_activation"))
        if out_act not in ("None", "Sigmoid"):
            raise LonerB200Error(f"unsupported output_activation {out_act}")
        self.n_input_dims, self.n_output_dims = n_input_dims, n_output_dims
        self.n_input_padded = _pad16(n_input_dims)
        if self.n_input_padded > 64 or n_output_dims > 16:
            raise LonerB200Error("Network: at most 64 inputs and 16 outputs")
        self.out_act = ops.OUT_ACT[out_act]
        self.dtype = torch.float32
        self.seed = seed
        w = int(network_config["n_neurons"])
        self.n_hidden = int(network_config["n_hidden_layers"])
        shapes = [(w, self.n_input_padded)] + [(w, w)] * (self.n_hidden - 1) + [(_pad16(n_output_dims), w)]
        self.shapes = shapes
        g = torch.Generator().manual_seed(seed)
        self.params = nn.Parameter(_init_mlp(shapes, g))

    def forward(self, x, out_act=None):
        x = x.float()
        if x.shape[1] < self.n_input_padded:
            x = torch.nn.functional.pad(x, (0, self.n_input_padded - x.shape[1]), value=1.0)
        return FusedNetwork.apply(x, self.params, self.n_hidden, self.n_output_dims,
                                  self.out_act if out_act is None else out_act)
