"""The library's own all-gather over peer memory (csrc/peer.cu, aether_peer_allgather) on ONE GPU: the `world` ranks are
emulated by `world` allocations on the same device, one stream per rank, so the store / flag / wait protocol runs for
real (the kernels of the ranks must overlap; they wait for each other's flags).  The multi-GPU run is
tools/peer_check.py under torchrun (bit-identical to the NCCL exchange; profiles/r02_peer_exchange.md)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world,block_bytes", [(2, 4096), (4, 12 * 1000), (8, 512 * 777), (3, 20)])
def test_peer_allgather_protocol_on_one_gpu(world, block_bytes):
    from aether_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    buf_off, flags_off = 256, 256 + (world * block_bytes + 255) // 256 * 256
    total = flags_off + 256
    copies = [torch.zeros(total, dtype=torch.uint8, device=dev) for _ in range(world)]
    states = [torch.zeros(4, dtype=torch.int32, device=dev) for _ in range(world)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(world)]
    base = (ctypes.c_void_p * world)(*[c.data_ptr() for c in copies])
    for epoch in range(1, 4):                      # three exchanges in a row over the same buffer
        want = torch.empty(world * block_bytes, dtype=torch.uint8)
        for r in range(world):
            blk = torch.randint(0, 256, (block_bytes,), dtype=torch.uint8)
            want[r * block_bytes:(r + 1) * block_bytes] = blk
            copies[r][buf_off + r * block_bytes: buf_off + (r + 1) * block_bytes] = blk.to(dev)
        torch.cuda.synchronize()
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                rc = lib.aether_peer_allgather(base, world, r, buf_off, block_bytes, flags_off, _lib.ptr(states[r]),
                                               ctypes.c_void_p(streams[r].cuda_stream))
                assert rc == 0
        torch.cuda.synchronize()
        if any(int(s[2]) for s in states):
            pytest.skip("the emulated ranks' kernels did not overlap on this device (a wait gave up)")
        for r in range(world):
            assert int(states[r][0]) == epoch and int(states[r][1]) == 0
            assert torch.equal(copies[r][buf_off:buf_off + world * block_bytes].cpu(), want), (r, epoch)
            flags = copies[r][flags_off:flags_off + 4 * world].view(torch.int32).cpu()
            assert all(int(flags[p]) == epoch for p in range(world) if p != r)


def test_peer_allgather_arguments():
    from aether_b200 import _lib
    lib = _lib.load()
    st = torch.zeros(4, dtype=torch.int32, device="cuda")
    b = torch.zeros(1024, dtype=torch.uint8, device="cuda")
    base = (ctypes.c_void_p * 1)(b.data_ptr())
    assert lib.aether_peer_allgather(base, 1, 0, 0, 512, 512, _lib.ptr(st), None) == 0      # one rank: nothing to do
    assert lib.aether_peer_allgather(base, 1, 1, 0, 512, 512, _lib.ptr(st), None) == -1
    assert lib.aether_peer_allgather(base, 17, 0, 0, 512, 512, _lib.ptr(st), None) == -1
    assert lib.aether_peer_allgather(base, 1, 0, 2, 512, 512, _lib.ptr(st), None) == -1     # unaligned offset
    assert lib.aether_peer_allgather(None, 1, 0, 0, 512, 512, _lib.ptr(st), None) == -1
