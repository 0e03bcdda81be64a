"""Pure-torch stand-in for the `tinycudann` module -- TEST INFRASTRUCTURE ONLY.

Lets the UNMODIFIED reference (`projects/neuralangelo/utils/modules.py:16,50,85`) run on the CPU in the
authoring container with a hash grid: `Encoding(3, config)` is an nn.Module with a flat fp32 `params`
whose forward is oracle/hashgrid_ref.py's arithmetic written with differentiable torch ops.  PARITY
UNPINNED with respect to the real tiny-cuda-nn (see hashgrid_ref.py); differences on purpose: fp32
features (tcnn returns fp16), no batch padding.
"""
import torch

from . import hashgrid_ref as hr


class Encoding(torch.nn.Module):
    def __init__(self, n_input_dims, encoding_config, seed=1337, dtype=None):
        super().__init__()
        assert n_input_dims == 3
        c = encoding_config
        self.tbl = hr.level_table(c["n_levels"], c["n_features_per_level"], c["log2_hashmap_size"],
                                  c["base_resolution"], c["per_level_scale"])
        self.n_input_dims = 3
        self.n_output_dims = c["n_levels"] * c["n_features_per_level"]
        self.params = torch.nn.Parameter(hr.init_table(self.tbl, seed=seed))

    def forward(self, x):
        F = self.tbl["n_features"]
        table = self.params.view(-1, F)
        x = x.detach().float()
        outs = []
        for lv in self.tbl["levels"]:
            cell, w = hr._locate(lv, x)
            idx = hr._corner_indices(lv, cell) + lv["offset"]
            ws = hr._weights(w)
            acc = 0
            for k in range(8):
                acc = acc + ws[k][:, None] * table[idx[k]]
            outs.append(acc)
        return torch.cat(outs, dim=-1)
