"""Generates tests/golden/ref_compress.npz with the REFERENCE's own quantiser: compress() imported from
/root/reference/inference/utils/compress_llama_weights.py (lines 15-58), run with the configuration its __main__ uses
(CompressionConfig(num_bits, group_size=32, group_dim=0, symmetric=False)) for 8 and 4 bits on seeded fp32 matrices.
Stores the inputs, the uint8 codes, the per-group min and scale (fp32, as torch computes them).
Run in the build container (needs /root/reference):  python tests/golden/make_compress_golden.py"""
import importlib.util
import os

import numpy as np
import torch

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_compress.npz")
SRC = "/root/reference/inference/utils/compress_llama_weights.py"


def main():
    spec = importlib.util.spec_from_file_location("ref_compress", SRC)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)  # defines CompressionConfig / compress / decompress (its __main__ block does not run)
    g = torch.Generator().manual_seed(0)
    out = {}
    for name, (o, i, scale) in {"a": (64, 48, 0.02), "b": (96, 40, 1.0), "c": (32, 8, 37.5)}.items():
        w = torch.randn(o, i, generator=g) * scale
        if name == "b":
            w[5, 3] = w[:32, 3].max()  # a tie for the maximum inside a group
            w[40:44, 7] = 0.0
        out["w_" + name] = w.numpy().copy()
        for bits in (8, 4):
            cfg = mod.CompressionConfig(num_bits=bits, group_size=32, group_dim=0, symmetric=False)
            data, mn, sc, shape = mod.compress(w.clone(), cfg)
            assert tuple(shape) == (o, i)
            out["q%d_%s" % (bits, name)] = data.reshape(o, i).numpy().copy()
            out["mn%d_%s" % (bits, name)] = mn.reshape(o // 32, i).numpy().copy()
            out["sc%d_%s" % (bits, name)] = sc.reshape(o // 32, i).numpy().copy()
    out["torch_version"] = np.array([torch.__version__])
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items() if k.startswith("q")})


if __name__ == "__main__":
    main()
