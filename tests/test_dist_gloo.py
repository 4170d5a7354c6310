"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: shard ranges, parameter broadcast,
pool gather order, max-over-ranks timing."""
import os
import sys
import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class _FakeLayer:
    def __init__(self, w, b): self.weights, self.biases = w, b


class _FakeModel:
    """Host-only stand-in with the attributes dist.broadcast_params touches (no CUDA here)."""
    def __init__(self, seed):
        rng = np.random.RandomState(seed)
        self.num_nets, self.separate_mean_var = 3, True
        self.layers = [_FakeLayer(rng.randn(3, 4, 5).astype(np.float32), rng.randn(3, 1, 5).astype(np.float32)),
                       _FakeLayer(rng.randn(3, 5, 2).astype(np.float32), rng.randn(3, 1, 2).astype(np.float32))]
        self.var_layers = [_FakeLayer(rng.randn(3, 5, 2).astype(np.float32), rng.randn(3, 1, 2).astype(np.float32))]
        self.mu, self.sigma = rng.randn(1, 4).astype(np.float32), rng.rand(1, 4).astype(np.float32) + 1
        self.max_lv, self.min_lv = rng.randn(1, 2).astype(np.float32), rng.randn(1, 2).astype(np.float32)
        if seed == 0:
            self._model_inds_list = [2, 0]
        self.device = torch.device('cpu')

    def _all_vars(self):
        v = [self.mu, self.sigma]
        for l in self.layers + self.var_layers:
            v += [l.weights, l.biases]
        return v + [self.max_lv, self.min_lv]

    def set_params(self, Ws, bs, vw, vb, mu, sigma, mx, mn):
        for l, w, b in zip(self.layers, Ws, bs):
            l.weights, l.biases = w, b
        self.var_layers[0].weights, self.var_layers[0].biases = vw, vb
        self.mu, self.sigma, self.max_lv, self.min_lv = mu, sigma, mx, mn

    @property
    def _model_inds(self): return self._model_inds_list

    @_model_inds.setter
    def _model_inds(self, v): self._model_inds_list = list(v)


class _FakePool:
    """CPU ring pool with the attributes dist.allgather_new_rows touches."""
    def __init__(self, cap, w=3):
        self._max_size, self._pointer, self.size = cap, 0, 0
        self.fields = {'observations': torch.zeros(cap, w), 'rewards': torch.zeros(cap, 1), 'terminals': torch.zeros(cap, 1, dtype=torch.bool)}

    def add_samples(self, s):
        n = len(s['rewards'])
        idx = (torch.arange(self._pointer, self._pointer + n) % self._max_size)
        for k in self.fields:
            self.fields[k][idx] = torch.as_tensor(s[k]).to(self.fields[k].dtype)
        self.advance_wrapped(n)

    def advance_wrapped(self, n):
        self._pointer = (self._pointer + n) % self._max_size
        self.size = min(self.size + n, self._max_size)

    def advance(self, n):
        assert self._pointer + n <= self._max_size
        self.advance_wrapped(n)


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    from mopo_b200 import dist as md
    r, w, _ = md.init_from_env(backend='gloo')
    assert (r, w) == (rank, world)
    model = _FakeModel(seed=rank)
    ref = _FakeModel(seed=0)
    md.broadcast_params(model, src=0)
    for a, b in zip(model._all_vars(), ref._all_vars()):
        assert np.array_equal(a, b)
    assert model._model_inds == [2, 0]
    lo, hi = md.shard_range(11, rank, world)
    n = hi - lo
    fields = {'observations': torch.arange(lo, hi, dtype=torch.float32)[:, None].repeat(1, 3),
              'terminals': (torch.arange(lo, hi) % 2 == 0)[:, None]}
    got = md.gather_pool_rows(fields, n, dst=0)
    ag = md.allgather_rows({'observations': torch.full((4, 3), float(rank)), 'terminals': torch.full((4, 1), rank == 1)})
    assert ag['observations'][:, 0].tolist() == [0.0] * 4 + [1.0] * 4
    assert ag['terminals'].dtype == torch.bool and ag['terminals'][:, 0].tolist() == [False] * 4 + [True] * 4
    # pool gather straight into the destination ring (in-place collective), then the wrapped fallback
    shard, full, want = _FakePool(10), _FakePool(17), _FakePool(17)
    for it in range(3):
        for r in range(world):
            rows = {'observations': torch.full((4, 3), float(10 * it + r)), 'rewards': torch.arange(4.0)[:, None] + 100 * r + it,
                    'terminals': torch.full((4, 1), (r + it) % 2 == 0)}
            if r == rank:
                shard.add_samples(rows)
            want.add_samples(rows)                       # what add_samples(concat over ranks) gives
        md.allgather_new_rows(shard, full, 4)
        assert full._pointer == want._pointer and full.size == want.size
        for k in full.fields:
            assert torch.equal(full.fields[k], want.fields[k]), (it, k)
    t = md.max_over_ranks(10.0 + rank, device='cpu')
    md.barrier()
    if rank == 0:
        assert got['observations'][:, 0].tolist() == list(range(11))
        assert got['terminals'].dtype == torch.bool and got['terminals'][:, 0].tolist() == [i % 2 == 0 for i in range(11)]
    else:
        assert got is None
    assert t == 10.0 + world - 1
    q.put((rank, lo, hi))
    dist.destroy_process_group()


def test_world2_gloo():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs: p.start()
    for p in procs: p.join(120)
    assert all(p.exitcode == 0 for p in procs)
    res = sorted(q.get() for _ in range(2))
    assert res == [(0, 0, 6), (1, 6, 11)]


def test_shard_ranges_cover_exactly():
    from mopo_b200.dist import shard_range
    for total in (0, 1, 7, 50000, 1000003):
        for world in (1, 2, 4, 8):
            rs = [shard_range(total, r, world) for r in range(world)]
            assert rs[0][0] == 0 and rs[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(rs, rs[1:]))
            sizes = [hi - lo for lo, hi in rs]
            assert max(sizes) - min(sizes) <= 1
