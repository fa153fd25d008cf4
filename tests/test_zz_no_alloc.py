"""The allocation-free entry point, b200c_emu_all_reduce_noalloc, against the allocating one.

Some callers of this library cannot allocate device memory at the moment they call it: a
self-test or health check that runs inside a serving process after the engine has handed every
free byte of HBM to its KV-cache pool (a cudaMalloc may fail, and it would dodge the engine's
memory accounting), or while another thread is capturing a CUDA graph (cudaMalloc is illegal
then).  For them the library keeps all scratch of the emulated all-reduce in one static device
array (about 3.6 MiB per GPU, messages up to 64 KiB per rank).

This test calls the entry point through the C ABI (ctypes), twice in a row with different
data, sizes and variants, and requires the same bits as b200c_emu_all_reduce (the all-reduce is
bit-exact everywhere in this suite).  It lives in a file that sorts last so that a run stopping
at the first failure has already run everything else; it must pass on its own in a fresh
process as well as after the other tests.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _call_noalloc(C, xs, outs, variant):
    rc = C.lib().b200c_emu_all_reduce_noalloc(
        len(xs), C.ptr_array([x.data_ptr() for x in xs]), C.ptr_array([o.data_ptr() for o in outs]),
        xs[0].numel(), {torch.bfloat16: C.BF16, torch.float16: C.F16, torch.float32: C.F32}[xs[0].dtype],
        variant, torch.cuda.current_stream().cuda_stream)
    return rc


def test_noalloc_matches_allocating_entry_point_twice_in_a_row():
    assert torch.cuda.is_available(), "-m gpu tests need a CUDA device"
    from vllm_distributed_b200 import _cabi as C
    from vllm_distributed_b200.emu import EmuGroup

    world = 4
    ref_group = EmuGroup(world, "cuda:0", max_bytes=1 << 20, ag_bytes=1 << 20, p2p_bytes=1 << 20)
    ref_group.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, 3000)
    gen = torch.Generator().manual_seed(2024)
    calls = [((3, 4096), torch.bfloat16, C.TWOSHOT),   # 24 KiB per rank
             ((2, 896), torch.bfloat16, C.ONESHOT),    # smaller than what the array last held
             ((1, 8192), torch.float32, C.AUTO),       # 32 KiB, fp32
             ((4, 8192), torch.bfloat16, C.AUTO)]      # exactly the 64 KiB limit
    # outputs are allocated up front: the calls under test must not need any allocation
    prepared = []
    for shape, dtype, variant in calls:
        xs = [torch.randn(*shape, generator=gen).to(dtype).cuda() for _ in range(world)]
        outs = [torch.empty_like(x) for x in xs]
        want = ref_group.all_reduce(xs, variant=variant)
        prepared.append((xs, outs, want, variant))
    torch.cuda.synchronize()

    # first call on this device: resolves the static array's address (outside any capture)
    xs, outs, want, variant = prepared[0]
    C.check(_call_noalloc(C, xs, outs, variant))
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.equal(outs[r].view(torch.uint8), want[r].view(torch.uint8)), f"call 0 rank {r}"

    free_before, _ = torch.cuda.mem_get_info()
    for i, (xs, outs, want, variant) in enumerate(prepared[1:], start=1):
        C.check(_call_noalloc(C, xs, outs, variant))   # back to back, same shared array
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before, "the allocation-free entry point changed free device memory"
    for i, (xs, outs, want, variant) in enumerate(prepared[1:], start=1):
        for r in range(world):
            assert torch.equal(outs[r].view(torch.uint8), want[r].view(torch.uint8)), f"call {i} rank {r}"

    # beyond the stated maximum: a clear error, nothing launched
    big = [torch.zeros(C.NOALLOC_MAX_BYTES // 2 + 8, dtype=torch.bfloat16, device="cuda") for _ in range(world)]
    rc = _call_noalloc(C, big, [torch.empty_like(b) for b in big], C.AUTO)
    assert rc == C.ETOOBIG
    assert b"no-allocation limit" in C.lib().b200c_last_error()
    # one-shot of 8 ranks x 32 KiB does not fit the receive region either
    eight = [torch.zeros(16 * 1024, dtype=torch.bfloat16, device="cuda") for _ in range(8)]
    assert _call_noalloc(C, eight, [torch.empty_like(b) for b in eight], C.ONESHOT) == C.ETOOBIG
    assert _call_noalloc(C, eight, [torch.empty_like(b) for b in eight], C.TWOSHOT) == C.OK
    torch.cuda.synchronize()
    ref_group.close()

    # three ranks at exactly the 64 KiB limit: 4096 vectors do not divide by 3, the two-shot
    # chunks are rounded up -- the static array carries that slack, so the documented limit holds
    ref3 = EmuGroup(3, "cuda:0", max_bytes=1 << 20, ag_bytes=1 << 20, p2p_bytes=1 << 20)
    ref3.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, 3000)
    xs3 = [torch.randn(C.NOALLOC_MAX_BYTES // 2, generator=gen).bfloat16().cuda() for _ in range(3)]
    outs3 = [torch.empty_like(x) for x in xs3]
    want3 = ref3.all_reduce(xs3, variant=C.TWOSHOT)
    for variant in (C.TWOSHOT, C.AUTO):
        C.check(_call_noalloc(C, xs3, outs3, variant))
        torch.cuda.synchronize()
        for r in range(3):
            assert torch.equal(outs3[r].view(torch.uint8), want3[r].view(torch.uint8)), f"3 ranks, rank {r}"
    ref3.close()
