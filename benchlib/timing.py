"""Timing protocol of bench.py: CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks; the K-step
loop is repeated R times so that every workload's timed region lasts at least `min_seconds` (a 1 ms region is noise: round 1's BENCH
and SCALE N=1 lines differed by 13 %), the MEDIAN repetition is reported; clocks and throttle reasons are sampled through NVML while
the loops run."""
import json
import os
import subprocess
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """SM clock, max clock, power and throttle reasons every ~20 ms (NVML in-process; nvidia-smi as a fall-back)."""

    BITS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, index):
        threading.Thread.__init__(self, daemon=True)
        self.index = index
        self.samples = []          # (sm_mhz, sm_max_mhz, reasons_bitmask)
        self.stop_flag = False
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(index))
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nvml = None

    @staticmethod
    def _physical_index(index):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                return int(vis.split(",")[index])
            except Exception:
                return index
        return index

    def _sample_smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5).stdout
        parts = [x.strip() for x in out.strip().split(",")]
        if len(parts) >= 6:
            mask = 0
            for (name, bit), v in zip(self.BITS, parts[2:6]):
                if v.lower().startswith("active"):
                    mask |= bit
            self.samples.append((float(parts[0]), float(parts[1]), mask))

    def run(self):
        while not self.stop_flag:
            try:
                if self.nvml is not None:
                    n = self.nvml
                    mhz = float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
                    mask = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
                    self.samples.append((mhz, self.max_mhz, mask))
                else:
                    self._sample_smi()
            except Exception:
                pass
            time.sleep(0.02)

    def mark(self):
        return len(self.samples)

    def summary(self, since=0, stop=False):
        if stop:
            self.stop_flag = True
            self.join(timeout=6)
        s = self.samples[since:]
        if not s:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no clock samples"], "samples": 0}
        sm = sorted(x[0] for x in s)
        reasons = [name for name, bit in self.BITS if any(x[2] & bit for x in s)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": s[0][1], "reasons": reasons, "samples": len(s)}


class Protocol(object):
    """barrier / timed loops / max over ranks for one process of the job."""

    def __init__(self, torch, dist, world, device, min_seconds):
        self.torch, self.dist, self.world, self.device, self.min_seconds = torch, dist, world, device, min_seconds

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, values):
        if self.world == 1:
            return [float(v) for v in values]
        t = self.torch.tensor(list(values), device=self.device, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    def timed(self, step, K, min_seconds=None):
        """R repetitions of `K` steps, each bracketed by barrier + synchronize; R chosen (identically on every rank) so that the
        repetitions last at least min_seconds in total.  Returns (median ms per K-step loop -- max over ranks per repetition --, R)."""
        torch = self.torch
        budget = self.min_seconds if min_seconds is None else min_seconds

        def one():
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            self.barrier()
            e0.record()
            for _ in range(K):
                step()
            e1.record()
            self.barrier()
            return e0.elapsed_time(e1)

        first = self.max_over_ranks([one()])[0]
        R = int(min(2000, max(3, -(-budget * 1e3 // max(first, 1e-3)))))
        times = [one() for _ in range(R)]
        times = self.max_over_ranks(times)             # max over ranks of every repetition, then the median repetition
        times.sort()
        return times[len(times) // 2], R
