"""Multi-GPU drivers: one process per GPU, torch.distributed (NCCL over NVLink) as plumbing.

The product path for more than one GPU is inside libjubann.so (csrc/comm.cu: jubann_comm_init,
collective jubann_query_*, jubann_mix_allgather / jubann_mix_apply, jubann_create_multi);
:func:`comm_init_from_torch` is the only piece of this module bench.py needs for it.  The
classes below compose the same steps from Python over torch.distributed: they carry the gloo
tests of the host-side logic on CPU and the device-pointer entry points' tests.

* :class:`ShardedEngine` -- rows sharded round-robin by global row index:
  every rank scans its shard (jubann_scan_local_dev), the shard-local top-k candidates are
  all-gathered, and jubann_merge_dev reproduces the single-table ranking bit for bit.
* :class:`ReplicaMixer`  -- MIX between replicas that each hold the whole table
  (mixable_versioned_table get_diff / put_diff, driven in the reference by
  jubatus/server/framework/mixer/linear_mixer.cpp:422-544): rows changed since the last
  exchange are packed into fixed-size records on the device, all-gathered, and applied
  everywhere; the host arbitrates versions (an existing row is overwritten only if its clock
  <= the incoming clock, the local clock moves past every clock seen).

Both classes talk to the engine through a small duck-typed surface, so the collective
plumbing and the version logic run under the gloo backend in the CPU tests.
"""
import numpy as np
import torch
import torch.distributed as dist


# ---- sharding arithmetic (include/jubann.h "Sharding") ---------------------------------
def shard_of(global_index, world):
    return global_index % world


def local_slot(global_index, world):
    return global_index // world


def global_index(local, rank, world):
    return local * world + rank


def local_count(global_rows, rank, world):
    return (global_rows - rank + world - 1) // world if global_rows > rank else 0


def gather_candidates(cand_key, cand_index, group=None):
    """[nq][k] per rank -> ([G][nq][k] keys, [G][nq][k] indices), the layout jubann_merge_dev takes."""
    world = dist.get_world_size(group)
    nq = cand_key.shape[0]
    # concatenation along dim 0 is the one output shape every backend accepts; viewed as [G][nq][k]
    all_k = torch.empty((world * nq,) + tuple(cand_key.shape[1:]), dtype=cand_key.dtype, device=cand_key.device)
    all_i = torch.empty((world * nq,) + tuple(cand_index.shape[1:]), dtype=cand_index.dtype, device=cand_index.device)
    dist.all_gather_into_tensor(all_k, cand_key.contiguous(), group=group)
    dist.all_gather_into_tensor(all_i, cand_index.contiguous(), group=group)
    all_k = all_k.view((world,) + tuple(cand_key.shape))
    all_i = all_i.view((world,) + tuple(cand_index.shape))
    return all_k, all_i


class ShardedEngine(object):
    """Row-sharded table over the ranks of a process group.

    `engine` is a jubatus_b200.Engine created with (shard_rank, shard_count) = (rank, world).
    Every rank must issue the same calls with the same arguments (queries are replicated).
    """

    def __init__(self, engine, group=None):
        self.engine = engine
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        assert engine.shard_rank == self.rank and engine.shard_count == self.world

    def set_rows_sig(self, row_index, sigs, norms=None):
        self.engine.set_rows_sig(row_index, sigs, norms)   # the library keeps rows with index % world == rank

    def query_sig_dev(self, d_qsigs, d_qnorms, k, flags=0):
        """d_qsigs: int64 [nq][w64] device tensor; returns (index int64, score float32, count int32) tensors."""
        e = self.engine
        nq = d_qsigs.shape[0]
        dev = d_qsigs.device
        stream = torch.cuda.current_stream(dev).cuda_stream if dev.type == "cuda" else 0
        ck = torch.empty((nq, k), dtype=torch.int32, device=dev)
        ci = torch.empty((nq, k), dtype=torch.int64, device=dev)
        qn = d_qnorms.data_ptr() if d_qnorms is not None else 0
        e.scan_local_dev(nq, d_qsigs.data_ptr(), qn, k, ck.data_ptr(), ci.data_ptr(), stream)
        all_k, all_i = gather_candidates(ck, ci, self.group)
        out_i = torch.zeros((nq, k), dtype=torch.int64, device=dev)
        out_s = torch.zeros((nq, k), dtype=torch.float32, device=dev)
        out_c = torch.zeros(nq, dtype=torch.int32, device=dev)
        e.merge_dev(nq, k, self.world, all_k.data_ptr(), all_i.data_ptr(), qn, flags, out_i.data_ptr(), out_s.data_ptr(),
                    out_c.data_ptr(), stream)
        return out_i, out_s, out_c


class ReplicaTable(object):
    """Host side of one replica's column_table: keys, (owner, clock) versions, key -> row index.

    `engine` stores the columns.  It needs: set_rows_sig(row_index, sigs, norms),
    diff_pack(row_index) -> uint8 tensor [n][record_bytes], diff_apply(records), record_bytes,
    and `device` (torch.device of its tensors).
    """

    def __init__(self, engine, my_id):
        self.engine = engine
        self.my_id = my_id
        self.keys = []
        self.versions = []      # (owner, clock)
        self.index = {}
        self.clock = 0
        self.mixed_clock = {}   # owner -> newest clock already exchanged (version_clock)

    def size(self):
        return len(self.keys)

    def add(self, keys, sigs, norms=None):
        """column_table::add for a batch: append or update in place, stamp (my_id, clock++)."""
        rows = np.zeros(len(keys), dtype=np.uint64)
        for i, key in enumerate(keys):
            at = self.index.get(key)
            if at is None:
                at = len(self.keys)
                self.keys.append(key)
                self.versions.append((self.my_id, self.clock))
                self.index[key] = at
            else:
                self.versions[at] = (self.my_id, self.clock)
            self.clock += 1
            rows[i] = at
        self.engine.set_rows_sig(rows, sigs, norms)
        return rows

    # mixable_versioned_table::get_diff: rows newer than what has been exchanged
    def get_diff(self):
        rows = [i for i, (owner, clock) in enumerate(self.versions)
                if owner not in self.mixed_clock or self.mixed_clock[owner] < clock]
        meta = [(self.keys[i], self.versions[i][0], self.versions[i][1]) for i in rows]
        return np.array(rows, dtype=np.uint64), meta

    def _note(self, owner, clock):
        if owner not in self.mixed_clock or self.mixed_clock[owner] < clock:
            self.mixed_clock[owner] = clock

    # mixable_versioned_table::put_diff over the gathered diffs
    def put_diff(self, metas, records, counts):
        """metas: per rank list of (key, owner, clock); records: uint8 [G][max_n][record_bytes]."""
        rb = self.engine.record_bytes
        take = []      # (src rank, src position, destination row)
        for r, meta in enumerate(metas):
            for j, (key, owner, clock) in enumerate(meta):
                at = self.index.get(key)
                store = True
                if at is None:
                    at = len(self.keys)
                    self.keys.append(key)
                    self.versions.append((owner, clock))
                    self.index[key] = at
                elif self.versions[at][1] <= clock:
                    self.versions[at] = (owner, clock)
                else:
                    store = False
                if self.clock <= clock:
                    self.clock = clock + 1
                self._note(owner, clock)
                if store:
                    take.append((r, j, at))
        if not take:
            return 0
        # two replicas may have shipped the same key in this round: the row is then named twice.
        # The scatter kernel writes one thread per word with no order between records, so only the
        # last record per destination row goes to the device (what applying them in order gives).
        last = {}
        for n_, (_, _, at) in enumerate(take):
            last[at] = n_
        take = [t for n_, t in enumerate(take) if last[t[2]] == n_]
        src = torch.tensor([r * records.shape[1] + j for r, j, _ in take], dtype=torch.int64, device=records.device)
        picked = records.reshape(-1, rb).index_select(0, src).contiguous()
        # records carry the SENDER's row index in their first 8 bytes; rewrite it to ours
        dst = torch.tensor([at for _, _, at in take], dtype=torch.int64, device=records.device)
        picked[:, :8] = dst.view(torch.uint8).reshape(-1, 8)
        self.engine.reserve(len(self.keys))
        self.engine.diff_apply(picked)
        return len(take)


class ReplicaMixer(object):
    """One MIX round = all-gather of every replica's diff (linear_mixer: gather get_diff from all,
    fold, broadcast put_diff -- here a single collective, broadcast_mixer.hpp:45-55 semantics)."""

    def __init__(self, table, group=None):
        self.table = table
        self.group = group
        self.world = dist.get_world_size(group)

    def mix(self):
        t = self.table
        rows, meta = t.get_diff()
        metas = [None] * self.world
        dist.all_gather_object(metas, meta, group=self.group)      # keys + versions: host metadata
        counts = [len(m) for m in metas]
        max_n = max(counts) if counts else 0
        if max_n == 0:
            return 0
        rb = t.engine.record_bytes
        mine = torch.zeros((max_n, rb), dtype=torch.uint8, device=t.engine.device)
        if len(rows):
            mine[: len(rows)] = t.engine.diff_pack(rows)
        gathered = torch.empty((self.world * max_n, rb), dtype=torch.uint8, device=t.engine.device)
        dist.all_gather_into_tensor(gathered, mine, group=self.group)   # row payloads: device records over NCCL
        gathered = gathered.view(self.world, max_n, rb)
        return t.put_diff(metas, gathered, counts)


def comm_init_from_torch(engine, group=None):
    """Joins `engine`'s handle to a communicator of its own inside libjubann.so (NCCL under the
    C-ABI): rank 0 asks the library for an id, torch.distributed only carries those 128 bytes to
    the other ranks -- the role ZooKeeper / the RPC layer has for the Jubatus servers.  After this
    the data path (sharded queries, MIX) runs in the library; no tensor goes through
    torch.distributed any more."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    box = [engine.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    engine.comm_init(box[0], rank, world)
    return rank, world


class EngineColumns(object):
    """Adapter from jubatus_b200.Engine (single-shard replica) to the surface ReplicaTable needs."""

    def __init__(self, engine):
        self.engine = engine
        self.device = torch.device("cuda", engine.device)
        self.record_bytes = engine.diff_record_bytes()
        self._rows = 0

    def set_rows_sig(self, rows, sigs, norms=None):
        self.engine.set_rows_sig(rows, sigs, norms)
        self._rows = self.engine.size()[0]

    def reserve(self, rows):
        """new rows arriving through MIX extend the table: append zero rows up to `rows`."""
        have = self.engine.size()[0]
        if rows > have:
            n = rows - have
            self.engine.set_rows_sig(np.arange(have, rows, dtype=np.uint64),
                                     np.zeros((n, self.engine.w64), dtype=np.uint64),
                                     np.zeros(n, dtype=np.float32) if self.engine.method == "euclid_lsh" else None)

    def diff_pack(self, rows):
        d_rows = torch.from_numpy(np.ascontiguousarray(rows).view(np.int64)).to(self.device)
        rec = torch.empty((len(rows), self.record_bytes), dtype=torch.uint8, device=self.device)
        self.engine.diff_pack_dev(len(rows), d_rows.data_ptr(), rec.data_ptr(),
                                  torch.cuda.current_stream(self.device).cuda_stream)
        return rec

    def diff_apply(self, records):
        self.engine.diff_apply_dev(records.shape[0], records.data_ptr(),
                                   torch.cuda.current_stream(self.device).cuda_stream)
        torch.cuda.current_stream(self.device).synchronize()
