"""Row-sharded data parallelism: one process per GPU, torch.distributed for plumbing.

SURVEY.md 8(e): rows are partitioned contiguously (shard r of G owns rows
[r*N/G, (r+1)*N/G)); centres / coefficients are replicated; the ONLY per-unit
exchange is one sum-all-reduce of the fp64 payload ([sums|counts|cost|n] per Lloyd
iteration, [grad|loss|weightSum] per LR pass).  This replaces Spark's
broadcast + reduceByKey / treeAggregate behind fit()
(reference call site: orangecontrib/spark/base/spark_ml_estimator.py:22).
Every rank then runs the identical finalize / optimizer step: no root, no broadcast.
"""


def bind_to_gpu_numa_node(device_index):
    """Pin this process to the CPU cores of the NUMA node its GPU hangs off (Linux sysfs), BEFORE it allocates pinned
    host memory: pages are then placed on that node (first touch) and every rank's host->device copies use its own
    memory controller and PCIe root instead of all ranks sharing node 0 -- which is what held the 4- and 8-GPU
    end-to-end numbers down (identical device times, 1.7x spread in wall time between runs).
    Returns {"node", "cpus"} or None when the topology cannot be read (single-node hosts, containers without sysfs).
    Sources, in order: sysfs (`numa_node` of the PCI device), then NVML's ideal-CPU mask for the device."""
    import os

    def _nvml_cpus():
        import pynvml
        import torch

        pynvml.nvmlInit()
        p = torch.cuda.get_device_properties(device_index)
        bdf = "%08x:%02x:%02x.0" % (getattr(p, "pci_domain_id", 0), p.pci_bus_id, p.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bdf.encode())
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        return [64 * i + b for i, w in enumerate(mask) for b in range(64) if (int(w) >> b) & 1]

    try:
        import torch

        p = torch.cuda.get_device_properties(device_index)
        bdf = "%04x:%02x:%02x.0" % (getattr(p, "pci_domain_id", 0), p.pci_bus_id, p.pci_device_id)
        node = -1
        try:
            with open("/sys/bus/pci/devices/%s/numa_node" % bdf) as f:
                node = int(f.read().strip())
        except OSError:
            pass
        cpus = set()
        if node >= 0:
            with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
                spec = f.read().strip()
            for part in spec.split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        else:
            cpus = set(_nvml_cpus())
            if not cpus or len(cpus) >= (os.cpu_count() or 1):
                return None  # no locality information: leave the affinity alone
        allowed = os.sched_getaffinity(0)
        cpus = sorted(cpus & allowed) or sorted(allowed)
        os.sched_setaffinity(0, cpus)
        return {"node": node, "cpus": len(cpus)}
    except Exception:
        return None


def shard_bounds(n_total, rank, world):
    """Rows [lo, hi) owned by `rank`."""
    return (n_total * rank) // world, (n_total * (rank + 1)) // world


class Comm:
    """Thin wrapper so the fit loops do not care whether a process group exists."""

    def __init__(self, group=None):
        import torch.distributed as dist

        self._dist = dist
        self.active = dist.is_available() and dist.is_initialized()
        self.group = group
        self.rank = dist.get_rank(group) if self.active else 0
        self.world = dist.get_world_size(group) if self.active else 1

    def all_reduce_sum(self, t):
        """In-place sum over ranks of a (device or host) tensor; no-op for world 1."""
        if self.active and self.world > 1:
            self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self.group)
        return t

    def all_reduce_max(self, t):
        if self.active and self.world > 1:
            self._dist.all_reduce(t, op=self._dist.ReduceOp.MAX, group=self.group)
        return t

    def broadcast(self, t, src=0):
        if self.active and self.world > 1:
            self._dist.broadcast(t, src=src, group=self.group)
        return t

    def any_rank(self, flag, device=None):
        """True on EVERY rank iff `flag` is true on at least one.  Error conditions found in one rank's shard (a bad
        label, a ragged column) are tested through this before raising: a rank that raises alone leaves the others
        blocked in the next collective."""
        if not (self.active and self.world > 1):
            return bool(flag)
        import torch

        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else "cpu"
        t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=device)
        self._dist.all_reduce(t, op=self._dist.ReduceOp.MAX, group=self.group)
        return bool(int(t.item()))

    def barrier(self):
        if self.active and self.world > 1:
            self._dist.barrier(group=self.group)
