"""Host-side placement for the end-to-end paths: pin a rank's threads to the CPU cores next to ITS GPU before it allocates
pinned host buffers (first touch then puts the buffers on that NUMA node), so that 8 ranks do not all stream their
host->device copies out of NUMA node 0 (e2e scaled 3.4x at 8 GPUs with every rank on node 0)."""
import os


def bind_to_gpu(local_rank):
    """Restricts this process to the CPUs NVML reports as local to GPU `local_rank`.  -> list of CPUs, or None when the
    affinity is unknown / cannot be set (the run then continues unbound)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(int(local_rank))
        n = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n + 63) // 64)
        cpus = [64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1 and 64 * i + b < n]
        allowed = set(os.sched_getaffinity(0))
        cpus = [c for c in cpus if c in allowed]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:  # noqa: BLE001
        return None
