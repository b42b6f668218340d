"""Two-rank check of the file-descriptor exchange used for the symmetric arenas (gloo, no GPU): every rank hands a
pipe's read end to every peer, writes its rank into the write end, and reads the peers' ranks through the received fds."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from b200dp.parallel.bootstrap import cleanup, setup_distributed  # noqa: E402
from b200dp.parallel.comm import _exchange_fds  # noqa: E402


def main() -> int:
    ctx = setup_distributed(use_cuda=False)
    r, w = os.pipe()
    got = _exchange_fds(ctx, [r], "test")
    os.write(w, bytes([ctx.rank + 65]))
    os.close(w)
    assert sorted(got) == [p for p in range(ctx.world_size) if p != ctx.rank], got
    import torch.distributed as dist
    dist.barrier()
    for peer, fds in got.items():
        assert len(fds) == 1
        assert os.read(fds[0], 1) == bytes([peer + 65])       # the peer's pipe, reached through the passed descriptor
        os.close(fds[0])
    empty = _exchange_fds(ctx, [], "none")                     # handing over nothing is legal (multicast set-up, non-root ranks)
    assert all(v == [] for v in empty.values())
    if ctx.rank == 0:
        print("FD EXCHANGE OK", flush=True)
    cleanup(ctx)
    return 0


if __name__ == "__main__":
    sys.exit(main())
