"""Multi-GPU plumbing for the rollout path: one process per GPU, rollout starts sharded
as contiguous row ranges, ensemble parameters replicated.  The data path has NO collective (rows are
independent, mopo.py:727-758); NCCL is used only (1) to broadcast the ensemble parameters after model
training and (2) to gather per-rank model_pool shards for the SAC sampler.  Works on gloo/CPU tensors
too, which is how the world_size-2 tests cover it without GPUs."""
import os
import numpy as np
import torch
import torch.distributed as dist


def init_from_env(backend=None):
    """Initialise torch.distributed from RANK / WORLD_SIZE / MASTER_* (torchrun). Returns (rank, world, local_rank)."""
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        os.environ.setdefault('MASTER_PORT', '29500')
        if backend is None:
            backend = 'nccl' if torch.cuda.is_available() else 'gloo'
        if backend == 'nccl':
            torch.cuda.set_device(local)
        # a long timeout: under torch.distributed only rank 0 trains the ensemble while the others wait in broadcast_params, and a
        # full-dataset BNN.train (max_epochs=None over 1-2 M rows) outlasts NCCL's default 10-minute collective timeout
        import datetime
        dist.init_process_group(backend=backend, rank=rank, world_size=world, timeout=datetime.timedelta(hours=6))
    return rank, world, local


def shard_range(total, rank, world):
    """Contiguous rows [lo, hi) of rank `rank`; the first `total % world` ranks get one extra row."""
    base, rem = divmod(int(total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def param_list(model):
    """Host parameter arrays in the .mat order (bnn.py:224-225)."""
    return model._all_vars()


def broadcast_params(model, src=0, device=None, sync_host=True):
    """Broadcast every ensemble parameter (and the elite list) from `src`: one flat buffer, one collective.

    NCCL + a device model (mopo_b200.models.bnn.BNN): the fp32 master images are packed device-to-device, broadcast, unpacked
    in place on the receivers and the tensor-core operand images are rebuilt there -- nothing crosses PCIe on the collective's
    path (sync_host=True then refreshes the host mirrors that save() reads; False leaves that to the caller).
    Otherwise (gloo / host models, the CPU tests): the same flat buffer built from the host arrays."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    if dist.get_backend() == 'nccl' and hasattr(model, '_device_images'):
        views = model._device_images()
        elites = torch.full((model.num_nets,), -1.0, dtype=torch.float32, device=views[0].device)
        if hasattr(model, '_model_inds_list'):
            elites[:len(model._model_inds_list)] = torch.tensor(model._model_inds_list, dtype=torch.float32)
        if dist.get_rank() == src:
            flat = torch.cat([v for v in views] + [elites])
        else:
            flat = torch.empty(sum(v.numel() for v in views) + model.num_nets, dtype=torch.float32, device=views[0].device)
        dist.broadcast(flat, src=src)
        if dist.get_rank() != src:
            off = 0
            for v in views:
                v.copy_(flat[off:off + v.numel()])
                off += v.numel()
            model._after_device_update(sync_host=sync_host)
            el = [int(e) for e in flat[off:].tolist() if e >= 0]
            if el:
                model._model_inds = el
        return
    arrs = param_list(model)
    flat = np.concatenate([np.asarray(a, np.float32).ravel() for a in arrs])
    elites = np.full((model.num_nets,), -1, np.float32)
    if hasattr(model, '_model_inds_list'):
        elites[:len(model._model_inds_list)] = model._model_inds_list
    buf = torch.from_numpy(np.concatenate([flat, elites]))
    if device is None:
        device = model.device if dist.get_backend() == 'nccl' else torch.device('cpu')
    buf = buf.to(device)
    dist.broadcast(buf, src=src)
    out = buf.cpu().numpy()
    off = 0
    new = []
    for a in arrs:
        n = int(np.prod(a.shape))
        new.append(out[off:off + n].reshape(a.shape).copy())
        off += n
    n_l = len(model.layers)
    Ws, bs = new[2:2 + 2 * n_l:2], new[3:3 + 2 * n_l:2]
    k = 2 + 2 * n_l
    vw = vb = None
    if model.separate_mean_var:
        vw, vb = new[k], new[k + 1]
        k += 2
    model.set_params(Ws, bs, vw, vb, new[0], new[1], new[k], new[k + 1])
    el = out[off:off + model.num_nets]
    el = [int(e) for e in el if e >= 0]
    if el:
        model._model_inds = el


def gather_pool_rows(fields, n_rows, dst=0):
    """Gather the first n_rows of every pool field to `dst`, rank-major (rank 0's rows first).
    fields: dict name -> tensor [>=n_rows, w].  Returns dict of concatenated tensors on dst, else None."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return {k: v[:n_rows] for k, v in fields.items()}
    world, rank = dist.get_world_size(), dist.get_rank()
    any_t = next(iter(fields.values()))
    counts = [torch.zeros(1, dtype=torch.int64, device=any_t.device) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([n_rows], dtype=torch.int64, device=any_t.device))
    counts = [int(c.item()) for c in counts]
    mx = max(counts)
    out = {}
    for k, v in fields.items():
        src = v[:n_rows]
        is_bool = src.dtype == torch.bool
        if is_bool:
            src = src.to(torch.uint8)
        pad = torch.zeros((mx,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
        pad[:n_rows] = src
        bufs = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(bufs, pad)
        if rank == dst:
            cat = torch.cat([b[:c] for b, c in zip(bufs, counts)], 0)
            out[k] = cat.bool() if is_bool else cat
    return out if rank == dst else None


def allgather_rows(fields):
    """All-gather equally sized row blocks from every rank, rank-major: fields = dict name -> tensor [n, w] (same n on every
    rank, e.g. the rows a never-terminating rollout just appended to this rank's model_pool shard).  Every rank receives the
    concatenation -- the SAC sampler's view of the whole rollout (BASELINE.json: "gather model_pool shards")."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return dict(fields)
    world = dist.get_world_size()
    out = {}
    for k, src in fields.items():
        is_bool = src.dtype == torch.bool
        src = (src.to(torch.uint8) if is_bool else src).contiguous()
        bufs = [torch.empty_like(src) for _ in range(world)]
        dist.all_gather(bufs, src)
        cat = torch.cat(bufs, 0)
        out[k] = cat.bool() if is_bool else cat
    return out


def allgather_new_rows(shard_pool, full_pool, n_new):
    """The pool gather of BASELINE.json: every rank's newest `n_new` rows (the rows its rollout shard just appended; the same
    count on every rank, as for never-terminating envs) land in EVERY rank's `full_pool`, rank-major, exactly where
    `full_pool.add_samples(concat over ranks)` would have put them -- but written by the collective itself: one
    all_gather_into_tensor per field straight into the field arrays at the ring offset (no staging concat, no host bounce),
    then the ring is advanced.  Falls back to a gathered index + add_samples when either span wraps around its ring."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    ptr, cap = shard_pool._pointer, shard_pool._max_size
    lo = (ptr - n_new) % cap
    src_contig = lo + n_new <= cap
    dptr, dcap = full_pool._pointer, full_pool._max_size
    if world == 1:
        rows = (torch.arange(lo, lo + n_new, device=shard_pool.fields['rewards'].device) % cap)
        full_pool.add_samples({k: v[rows] for k, v in shard_pool.fields.items()})
        return
    if src_contig and dptr + world * n_new <= dcap and hasattr(full_pool, 'advance'):
        for k, v in shard_pool.fields.items():
            src = v[lo:lo + n_new]
            dst = full_pool.fields[k][dptr:dptr + world * n_new]
            if src.dtype == torch.bool:
                src, dst = src.view(torch.uint8), dst.view(torch.uint8)
            dist.all_gather_into_tensor(dst, src)
        full_pool.advance(world * n_new)
        return
    rows = (torch.arange(lo, lo + n_new, device=shard_pool.fields['rewards'].device) % cap)
    full_pool.add_samples(allgather_rows({k: v[rows] for k, v in shard_pool.fields.items()}))


def max_over_ranks(value, device=None):
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    dev = device if device is not None else ('cuda' if dist.get_backend() == 'nccl' else 'cpu')
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier():
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


def shutdown():
    """Tear the process group down (quiet exit under torchrun)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
