"""
ms_diffuse_molecules_static: diffusion for callers that cannot allocate device memory.

Some hosts that embed this library must not allocate on the device when they call into it:
their device memory is owned by a pool they set up at start, or they cannot afford the
device-wide synchronisation that cudaMalloc / cudaFree cause in the middle of a step loop.
For them the library carries its own fixed scratch (a 16 MiB `static __device__` array) and
the entry point takes no scratch argument.  This test drives it through the C ABI only and
checks, on two calls in a row (the second on a smaller map, so stale contents of the shared
array from the first call are still there), that its result is bit-identical to
ms_diffuse_molecules with caller scratch; and that maps beyond the built-in limit are
refused instead of overrunning the array.

Collected last on purpose (file name): if it fails under `-x`, everything else has run.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _weights(nat, diffs, dev):
    from magicsoup_b200.world import diffusion_weights

    a = torch.tensor([diffusion_weights(d)[0] for d in diffs], dtype=torch.float32, device=dev)
    b = torch.tensor([diffusion_weights(d)[1] for d in diffs], dtype=torch.float32, device=dev)
    return a, b


def _with_caller_scratch(nat, mm, a, b):
    m, H, W = mm.shape
    n = int(nat.lib().ms_diffuse_scratch_bytes(m, H, W))
    scratch = torch.empty(n, dtype=torch.uint8, device=mm.device)
    out = mm.clone()
    nat.check(nat.lib().ms_diffuse_molecules(nat.ptr(out), nat.ptr(a), nat.ptr(b), m, H, W, nat.ptr(scratch), n,
                                             nat.stream_ptr(mm.device)))
    return out


def test_static_scratch_diffusion_matches_and_rejects_oversize():
    from magicsoup_b200 import _native as nat

    nat.require_cuda()
    dev = torch.device("cuda", torch.cuda.current_device())
    lib = nat.lib()
    limit = int(lib.ms_diffuse_static_max_floats())
    assert limit == 4 * 1024 * 1024

    # two calls in a row: a tiled-path map that fills most of the built-in array, then a small
    # generic-path map whose scratch region still holds the first call's values
    cases = [((12, 512, 640), 1), ((3, 33, 47), 2)]
    for (m, H, W), seed in cases:
        assert m * H * W <= limit
        g = torch.Generator(device="cpu").manual_seed(seed)
        mm = (torch.randn(m, H, W, generator=g) + 10.0).abs().to(dev)
        a, b = _weights(nat, [1.0, 0.1, 0.01, 0.0, 0.5][:m] + [0.3] * max(0, m - 5), dev)
        expected = _with_caller_scratch(nat, mm, a, b)
        got = mm.clone()
        for step in range(2):  # and twice on the same map: the sums must be cleared every call
            nat.check(lib.ms_diffuse_molecules_static(nat.ptr(got), nat.ptr(a), nat.ptr(b), m, H, W,
                                                      nat.stream_ptr(dev)))
            torch.cuda.synchronize()
            assert torch.equal(got, expected), f"map {m}x{H}x{W}, call {step}"
            expected = _with_caller_scratch(nat, expected, a, b)

    # beyond the limit: refused with a clear error, nothing written
    m, H, W = 5, 1024, 1024
    assert m * H * W > limit
    big = torch.ones(m, H, W, device=dev)
    a, b = _weights(nat, [0.5] * m, dev)
    rc = lib.ms_diffuse_molecules_static(nat.ptr(big), nat.ptr(a), nat.ptr(b), m, H, W, nat.stream_ptr(dev))
    assert rc == -5
    msg = lib.ms_diffuse_static_last_error().decode()
    assert "exceeds the built-in scratch" in msg and str(limit) in msg
    with pytest.raises(nat.NativeError, match="beyond the supported maximum"):
        nat.check(rc)
    torch.cuda.synchronize()
    assert bool((big == 1.0).all())
