"""One lattice over several GPUs: row slabs with halo rows written over NVLink.

No counterpart in the reference (it has no domain decomposition).  One process per GPU, started with
``torchrun``; process g owns rows ``[g*lenX/G, (g+1)*lenX/G)`` of the lattice.  ``torch.distributed`` is
used once, to hand each process the CUDA IPC handles of its two neighbours' buffers.  After that a call of two
or more sweeps keeps every slab in the shared memory of its GPU for the whole call and hands the first / last row
of every colour pass to the neighbouring GPUs as packets stored into their memory over NVLink (one launch per GPU);
a call of a single sweep stores the boundary rows into the neighbours' halo rows from inside the tiled sweep kernel,
which also carries the neighbour barrier -- no NCCL call and no host involvement per sweep either way.
The random stream is indexed by whole-lattice rows, so the result is bit-identical to the single-GPU run.
"""
import os

import numpy as np

from . import _cabi


class SlabLattice:
    def __init__(self, model_code, lenX, lenY, seed, dist=None, device=None):
        self.dist = dist if (dist is not None and dist.is_initialized()) else None
        self.rank = self.dist.get_rank() if self.dist else 0
        self.world = self.dist.get_world_size() if self.dist else 1
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", "0")) if self.world > 1 else 0
        self.lenX, self.lenY = lenX, lenY
        self.rows = lenX // self.world
        self.lat = _cabi.Lattices(model_code, lenX, lenY, seed=seed, layout=_cabi.LAYOUT_INT8, device=device,
                                  slab=(self.rank, self.world))
        if self.world > 1:
            mine = self.lat.slab_ipc_export()
            everyone = [None] * self.world
            self.dist.all_gather_object(everyone, mine)
            self.lat.slab_connect(everyone[(self.rank - 1) % self.world], everyone[(self.rank + 1) % self.world])
            self.dist.barrier()

    def init_random(self):
        self.lat.init_random()

    def set_accept_tables(self, thr):
        self.lat.set_accept_tables(thr)

    def upload_rows(self, rows_int8):
        """This process's rows [rank*rows, (rank+1)*rows) of the lattice; collective (refreshes the halos)."""
        self.lat.upload(0, rows_int8)
        if self.dist:
            self.dist.barrier()  # every slab has its rows before anyone pushes halos into them
        self.lat.slab_push_halos()

    def sweep(self, n=1):
        self.lat.sweep(n)

    def check(self):
        if self.lat.slab_status():
            raise RuntimeError("a neighbouring slab did not reach the barrier in time; results are invalid")

    def download_rows(self):
        return self.lat.download(0)

    def observables(self):
        """(bond sum, second observable) of the whole lattice (sum over slabs)."""
        b, s = self.lat.tracked()
        v = np.array([float(b[0]), float(s[0])])
        if self.dist:
            import torch
            t = torch.from_numpy(v).cuda()
            self.dist.all_reduce(t)
            v = t.cpu().numpy()
        return int(v[0]), int(v[1])

    def gather(self):
        """The whole lattice on every process (for checks; int8 [lenX, lenY])."""
        mine = self.download_rows()
        if not self.dist:
            return mine
        import torch
        t = torch.from_numpy(mine).cuda()
        out = torch.empty((self.world,) + tuple(mine.shape), dtype=torch.int8, device="cuda")
        self.dist.all_gather_into_tensor(out, t)
        return out.cpu().numpy().reshape(self.lenX, self.lenY)
