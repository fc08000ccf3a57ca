"""Generate tests/golden/ref_producers.npz by RUNNING THE REFERENCE's RMSNorm and SiluAndMul
modules (nanovllm/layers/layernorm.py, nanovllm/layers/activation.py, imported unmodified from
/root/reference) on CPU in this build container:

    python tests/golden/make_golden_producers.py

The modules decorate their forwards with ``@torch.compile``; inductor's C++ backend does not
build in this container (CppCompileError), so dynamo is disabled and the SAME Python source
runs eagerly (every torch op rounds to its output dtype).  The reference's GPU run executes
the inductor-fused version, which skips the intermediate bf16 roundings -- <= 1 bf16 ulp away;
that is why the CUDA parity bar for these producers is a stated tolerance, not bit equality.

Stored as bit patterns (bf16 -> uint16) with the inputs alongside.
"""
import os
import sys

os.environ["TORCHDYNAMO_DISABLE"] = "1"
os.environ.setdefault("CUDA_VISIBLE_DEVICES", "")
REF = "/root/reference"
sys.path.insert(0, REF)

import numpy as np  # noqa: E402
import torch  # noqa: E402
from nanovllm.layers.activation import SiluAndMul  # noqa: E402
from nanovllm.layers.layernorm import RMSNorm  # noqa: E402
import nanovllm.layers.layernorm as _ln  # noqa: E402

assert _ln.__file__.startswith(REF), _ln.__file__


def bits(t):
    return t.detach().contiguous().view(torch.int16).numpy().view(np.uint16).copy()


def main():
    out = {}
    g = torch.Generator().manual_seed(20261105)
    for tag, (M, K) in {"a": (7, 1024), "b": (3, 4096), "c": (5, 136)}.items():
        x = (torch.randn(M, K, generator=g) * 1.7).to(torch.bfloat16)
        x[1] *= 1e-3
        if M > 4:
            x[4].zero_()
        r = torch.randn(M, K, generator=g).to(torch.bfloat16)
        w = (1.0 + 0.3 * torch.randn(K, generator=g)).to(torch.bfloat16)
        norm = RMSNorm(K, eps=1e-6).to(torch.bfloat16)
        norm.weight.data.copy_(w)
        with torch.no_grad():
            h0 = norm(x.clone())
            h1, res1 = norm(x.clone(), r.clone())
        out[f"rms_{tag}_x"], out[f"rms_{tag}_r"], out[f"rms_{tag}_w"] = bits(x), bits(r), bits(w)
        out[f"rms_{tag}_h"], out[f"rms_{tag}_hres"], out[f"rms_{tag}_res"] = bits(h0), bits(h1), bits(res1)
    act = SiluAndMul()
    for tag, (M, I) in {"a": (6, 3072), "b": (2, 12288), "c": (9, 200)}.items():
        gu = (torch.randn(M, 2 * I, generator=g) * 2.5).to(torch.bfloat16)
        gu[0, :8] = torch.tensor([0.0, -0.0, 30.0, -30.0, 88.0, -88.0, 1e-3, -1e-3]).to(torch.bfloat16)
        with torch.no_grad():
            h = act(gu.clone())
        out[f"silu_{tag}_x"], out[f"silu_{tag}_h"] = bits(gu), bits(h)
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_producers.npz")
    np.savez_compressed(dst, **out)
    print(f"wrote {dst}: {len(out)} arrays, {os.path.getsize(dst)} bytes")


if __name__ == "__main__":
    main()
