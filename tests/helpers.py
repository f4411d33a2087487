import hashlib

import numpy as np
import torch


def sd_sha256(sd):
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(sd[k].contiguous().numpy().tobytes())
    return h.hexdigest()


def lr_cases(npz):
    names = sorted({k.split(".")[0] for k in npz.files if k.endswith(".dur") and not k.startswith("round")})
    for name in names:
        dur = npz[f"{name}.dur"]
        ml = int(npz[f"{name}.max_len"])
        ml = None if ml < 0 else ml
        if f"{name}.x" in npz.files:
            x = torch.from_numpy(npz[f"{name}.x"])
        else:
            x = torch.randn(64, 32, 256, generator=torch.Generator().manual_seed(3))
        if bool(npz[f"{name}.x_is_bf16"]):
            x = x.to(torch.bfloat16)
        yield name, x, dur, ml


def check_lr_output(npz, name, out: torch.Tensor, mel_len):
    """out: torch tensor (cpu) of the case's dtype; bit-exact comparison with the fixture."""
    np.testing.assert_array_equal(np.asarray(mel_len), npz[f"{name}.mel_len"])
    assert np.asarray(mel_len).dtype == np.int64
    assert (out.dtype == torch.bfloat16) == bool(npz[f"{name}.out_is_bf16"])
    if f"{name}.out" in npz.files:
        exp = npz[f"{name}.out"]
        assert tuple(out.shape) == exp.shape, (name, out.shape, exp.shape)
        got = out.float().numpy()
        assert got.tobytes() == exp.tobytes(), name
    else:
        assert list(out.shape) == npz[f"{name}.out_shape"].tolist()
        sha = hashlib.sha256(out.contiguous().numpy().tobytes()).hexdigest()
        assert sha == str(npz[f"{name}.out_sha256"]), name
