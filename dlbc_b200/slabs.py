"""Host-side logic of the z-slab decomposition (one process per GPU).

Nothing here computes lattice physics; it is the bookkeeping a host program needs around the C ABI when
nranks > 1, kept separate from `lattice.py` so that it can be tested on CPU with the gloo backend
(tests/test_multi_rank_gloo.py):

* which global z-planes a rank owns (`slab_range`: the reference's M.nc = [1, 1, nranks],
  src/dlbc/parallel.d:190-237);
* which populations cross a slab face (`UP_Q` / `DOWN_Q`, from connectivity.d:401-418) -- the library sends
  only those 5 planes per component and side instead of the reference's 2 layers x 19
  (fields/parallel.d:58-103);
* broadcasting rank 0's NCCL unique id and reducing timings over ranks with torch.distributed
  (plumbing only: the data path uses NCCL send/recv inside libdlbc_cuda.so).
"""
from __future__ import annotations

import numpy as np

from .connectivity import d3q19
from .lattice import slab_range  # noqa: F401  (re-exported)

#: populations that leave a slab through its +z face (c_z = +1): they fill the upper neighbour's ghost plane 0
UP_Q = tuple(int(q) for q in np.nonzero(d3q19.velocities[:, 2] == 1)[0])
#: populations that leave through the -z face (c_z = -1): they fill the lower neighbour's ghost plane nz+1
DOWN_Q = tuple(int(q) for q in np.nonzero(d3q19.velocities[:, 2] == -1)[0])


def neighbours(rank: int, nranks: int):
    """(down, up) ranks of the periodic ring."""
    return (rank - 1) % nranks, (rank + 1) % nranks


def broadcast_unique_id(dist, rank: int, make_id, nbytes: int, device="cpu") -> bytes:
    """Rank 0 calls make_id() -> bytes; every rank returns the same bytes."""
    import torch

    t = torch.zeros(nbytes, dtype=torch.uint8, device=device)
    if rank == 0:
        raw = make_id()
        if len(raw) != nbytes:
            raise ValueError(f"unique id has {len(raw)} bytes, expected {nbytes}")
        t.copy_(torch.frombuffer(bytearray(raw), dtype=torch.uint8))
    dist.broadcast(t, 0)
    return bytes(t.cpu().numpy().tobytes())


def max_over_ranks(dist, x: float, device="cpu") -> float:
    """Timing rule: a multi-GPU number is the max over ranks."""
    if dist is None:
        return float(x)
    import torch

    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def ring_exchange(dist, rank: int, nranks: int, send_up, send_down):
    """Send `send_up` to rank+1 and `send_down` to rank-1 (periodic); returns (from_down, from_up).
    CPU/gloo helper used by the tests to emulate the exchange; torch tensors in, torch tensors out."""
    import torch

    down, up = neighbours(rank, nranks)
    from_down = torch.empty_like(send_up)
    from_up = torch.empty_like(send_down)
    if nranks == 1:
        return send_up.clone(), send_down.clone()
    reqs = [dist.isend(send_up.contiguous(), up, tag=1), dist.isend(send_down.contiguous(), down, tag=2),
            dist.irecv(from_down, down, tag=1), dist.irecv(from_up, up, tag=2)]
    for r in reqs:
        r.wait()
    return from_down, from_up
