"""Host-side plumbing for the two multi-GPU workloads.  One process per GPU.

* robots partition over ranks with no data-path communication (`partition_robots`);
* one large controller splits its samples over ranks (`shard_slice` mirrors the slice `qpmpc_create` assigns) and needs
  one NCCL all-gather of a small record per tick; the ncclUniqueId is handed out with `broadcast_bytes` over whatever
  `torch.distributed` backend the launcher initialised (gloo on CPU hosts, nccl on GPU boxes).
"""
from __future__ import annotations

from .batched import partition_robots  # noqa: F401  (re-exported)


def shard_slice(num_samples: int, rank: int, world: int):
    """Contiguous slice [begin, end) of the global sample indices owned by `rank` (same rule as csrc/qpmpc_api.cu)."""
    per = (num_samples + world - 1) // world
    begin = min(per * rank, num_samples)
    return begin, min(begin + per, num_samples)


def broadcast_bytes(payload: bytes | None, nbytes: int, src: int = 0) -> bytes:
    """Broadcast `nbytes` bytes from rank `src` over the initialised torch.distributed group."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        assert payload is not None
        return payload
    dev = 'cuda' if dist.get_backend() == 'nccl' else 'cpu'
    t = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
    if dist.get_rank() == src:
        t.copy_(torch.frombuffer(bytearray(payload), dtype=torch.uint8))
    dist.broadcast(t, src)
    return bytes(t.cpu().numpy().tobytes())


def all_gather_bytes(payload: bytes) -> list:
    """Every rank contributes `payload` (same length everywhere); returns the payloads in rank order."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return [payload]
    dev = 'cuda' if dist.get_backend() == 'nccl' else 'cpu'
    mine = torch.frombuffer(bytearray(payload), dtype=torch.uint8).to(dev)
    parts = [torch.zeros_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(parts, mine)
    return [bytes(p.cpu().numpy().tobytes()) for p in parts]


def make_sharded_controller(cfg, rank: int | None = None, world: int | None = None, exchange: str = 'peer'):
    """`Sampling_MPC` owning shard `rank` of `world` (ranks from torch.distributed).
    exchange = 'peer': the record exchange is fused into the tick kernel over NVLink peer memory (2..8 ranks of one
    node, one launch per tick); 'nccl': partial kernel + ncclAllGather + finish kernel."""
    import torch.distributed as dist
    from .sampling_mpc import Sampling_MPC
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    if exchange not in ('peer', 'nccl'):
        raise ValueError(f'unknown exchange {exchange!r}')
    ctrl = Sampling_MPC(cfg, shard_rank=rank, shard_world=world)
    if world > 1 and exchange == 'peer':
        ctrl.peer_attach(all_gather_bytes(ctrl.peer_export()))
    elif world > 1:
        uid = Sampling_MPC.nccl_unique_id() if rank == 0 else None
        ctrl.init_nccl(broadcast_bytes(uid, 128, 0))
    return ctrl
