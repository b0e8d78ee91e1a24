"""Page-locked host arrays placed next to a GPU.

The host-buffer entry (ops.HostPipeline, pdec_burgers_pair_loss_host) moves 2 x 69 MB per step over PCIe; on a two-socket box a buffer
that lives on the other socket's memory crosses the inter-socket link in both directions and, with eight ranks doing the same, that
link and the remote memory controller become the bound.  Linux places a page on the NUMA node of the thread that first touches it, so
the allocation is done with the calling thread restricted to the cores NVML reports as local to the GPU, then the old affinity is
restored.  (The reference has no counterpart: it feeds pageable NumPy arrays through session.run.)"""
import os

import torch


def gpu_local_cpus(device_index):
    """Set of CPU ids local to CUDA device `device_index` (NVML ideal CPU affinity), or None when NVML does not answer."""
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            props = torch.cuda.get_device_properties(device_index)
            handle = None
            uuid = getattr(props, "uuid", None)
            if uuid is not None:
                try:
                    handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode())
                except Exception:
                    handle = None
            if handle is None:
                handle = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            ncpu = os.cpu_count() or 1
            words = pynvml.nvmlDeviceGetCpuAffinity(handle, (ncpu + 63) // 64)
            cpus = {w * 64 + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
            return cpus or None
        finally:
            pynvml.nvmlShutdown()
    except Exception:
        return None


def pinned_empty(shape, device=None, dtype=torch.float32):
    """Page-locked CPU tensor whose pages are first touched on the NUMA node of `device` (default: the current CUDA device)."""
    index = torch.cuda.current_device() if device is None else torch.device(device).index or 0
    cpus = None if os.environ.get("PDEC_NO_NUMA") else gpu_local_cpus(index)
    old = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else None
    try:
        if cpus and old is not None:
            allowed = cpus & old or cpus
            try:
                os.sched_setaffinity(0, allowed)
            except OSError:
                pass
        t = torch.empty(shape, dtype=dtype).pin_memory()
        t.zero_()
        return t
    finally:
        if old is not None:
            try:
                os.sched_setaffinity(0, old)
            except OSError:
                pass
