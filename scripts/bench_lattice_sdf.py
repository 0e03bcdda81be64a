"""Block query of extract_mesh (129^3 lattice, 2048^3 spacing): one-kernel na_lattice_sdf against encode planes + SDF-only MLP.
Usage: python scripts/bench_lattice_sdf.py [blocks]"""
import json
import sys

import torch

sys.path.insert(0, ".")
from neuralangelo_b200.config import default_model_config  # noqa: E402
from neuralangelo_b200.modules import NeuralSDF  # noqa: E402


def main():
    blocks = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    cfg = default_model_config()
    net = NeuralSDF(cfg.object.sdf).to(dev).eval()
    with torch.no_grad():
        net.tcnn_encoding.params.uniform_(-1e-4, 1e-4)
    net.warm_up_end = 0
    net.set_active_levels(current_iter=10 ** 9)
    step = 2.0 / 2048
    ar = torch.arange(129, device=dev, dtype=torch.float32) * step
    origins = [(-1.0 + 128 * step * (b % 16), -1.0 + 128 * step * ((b // 16) % 16), -1.0 + 128 * step * (5 + b // 256))
               for b in range(blocks)]
    out = {}
    with torch.no_grad():
        for name, fused in (("two_launch", False), ("one_kernel", True)):
            for o in origins[:2]:
                net.sdf_lattice(ar + o[0], ar + o[1], ar + o[2], fused=fused)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for o in origins:
                net.sdf_lattice(ar + o[0], ar + o[1], ar + o[2], fused=fused)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / blocks
            out[name] = {"ms_per_block": ms, "Mpoints_per_s": 129 ** 3 / ms / 1e3}
        a = net.sdf_lattice(ar + origins[3][0], ar + origins[3][1], ar + origins[3][2], fused=False)
        b = net.sdf_lattice(ar + origins[3][0], ar + origins[3][1], ar + origins[3][2], fused=True)
        out["max_abs_diff"] = float((a - b).abs().max())
        out["max_abs"] = float(a.abs().max())
        from neuralangelo_b200 import _lib
        import ctypes
        so = ctypes.CDLL(_lib.lib()._name)
        if hasattr(so, "na_lattice_dev_set"):      # developer build (-DNA_LATTICE_DEV): which role bounds the tile?
            for mode in (1, 2, 3, 4, 8, 10, 19, 26):
                so.na_lattice_dev_set(mode)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for o in origins:
                    net.sdf_lattice(ar + o[0], ar + o[1], ar + o[2], fused=True)
                e1.record()
                torch.cuda.synchronize()
                out[f"dev_mode_{mode}"] = e0.elapsed_time(e1) / blocks
            so.na_lattice_dev_set(0)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
