"""MemoryManager / MemoryPressure -- same bookkeeping surface as the reference's
src/memory_manager.py (:14-304): pressure levels with hysteresis, degradation recommendations,
OOM counters, `get_memory_stats()` keys.

Differences that follow from the B200 build: libtcd allocates with cudaMalloc into static arenas,
outside torch's caching allocator, so usage is read from the driver (cudaMemGetInfo via
torch.cuda.mem_get_info) instead of torch's reserved bytes, and `clear_cache` has nothing to
release.  north_star rules out a CPU path, so `cpu_fallback` is never recommended.
"""
import logging
import time
from enum import Enum
from typing import Any, Dict

logger = logging.getLogger(__name__)


class MemoryPressure(Enum):
    NORMAL = "normal"      # < 75 %
    HIGH = "high"          # 75-85 %
    CRITICAL = "critical"  # 85-95 %
    EXTREME = "extreme"    # > 95 %


def _cuda():
    try:
        import torch
        return torch if torch.cuda.is_available() else None
    except Exception:
        return None


class MemoryManager:
    def __init__(self, device: str = "cuda:0", warning_threshold: float = 0.75, high_threshold: float = 0.85,
                 critical_threshold: float = 0.95, check_interval: float = 1.0, hysteresis: float = 0.05):
        self.device = device
        self.warning_threshold = warning_threshold
        self.high_threshold = high_threshold
        self.critical_threshold = critical_threshold
        self.check_interval = check_interval
        self.hysteresis = hysteresis
        self.current_pressure = MemoryPressure.NORMAL
        self.last_check_time = 0.0
        self.last_warning_time = 0.0
        self.warning_cooldown = 30.0
        self.oom_events = 0
        self.recoveries = 0
        self.degradation_level = 0
        self.cuda_available = _cuda() is not None
        if not self.cuda_available:
            logger.warning("CUDA not available - Memory Manager idle")

    def _usage(self):
        """-> (used_bytes, total_bytes) of the whole device."""
        free, total = _cuda().cuda.mem_get_info(self.device)
        return total - free, total

    def _classify(self, frac: float) -> MemoryPressure:
        if frac >= self.critical_threshold:
            return MemoryPressure.EXTREME
        if frac >= self.high_threshold:
            return MemoryPressure.CRITICAL
        if frac >= self.warning_threshold:
            return MemoryPressure.HIGH
        # below the warning line: leave a raised state only once usage clears the hysteresis band
        if self.current_pressure != MemoryPressure.NORMAL and frac >= self.warning_threshold - self.hysteresis:
            return self.current_pressure
        return MemoryPressure.NORMAL

    def check_memory_pressure(self, force: bool = False) -> MemoryPressure:
        if not self.cuda_available:
            return MemoryPressure.NORMAL
        now = time.time()
        if not force and now - self.last_check_time < self.check_interval:
            return self.current_pressure
        self.last_check_time = now
        try:
            used, total = self._usage()
            frac = used / total if total > 0 else 0.0
            new = self._classify(frac)
            if new != self.current_pressure:
                logger.warning("GPU memory pressure changed: %s -> %s (%.1f%% of %.1f GB in use)",
                               self.current_pressure.value, new.value, frac * 100, total / 1e9)
                self.current_pressure = new
            if new != MemoryPressure.NORMAL and now - self.last_warning_time >= self.warning_cooldown:
                logger.warning("Sustained GPU memory pressure: %s (%.1f%% used, degradation level %d)", new.value,
                               frac * 100, self.degradation_level)
                self.last_warning_time = now
            return new
        except Exception as e:
            logger.error("Failed to check GPU memory pressure: %s", e)
            return MemoryPressure.NORMAL

    def reduce_memory_usage(self, level: MemoryPressure) -> Dict[str, Any]:
        rec = {"clear_cache": False, "reduce_batch_size": False, "disable_frame_buffering": False,
               "reduce_input_size": False, "suggested_input_size": None, "suggested_batch_size": None}
        if level in (MemoryPressure.HIGH, MemoryPressure.CRITICAL, MemoryPressure.EXTREME):
            rec["clear_cache"] = True
        if level in (MemoryPressure.CRITICAL, MemoryPressure.EXTREME):
            rec["reduce_batch_size"] = True
            rec["suggested_batch_size"] = 1
        if level == MemoryPressure.EXTREME:
            rec["disable_frame_buffering"] = True
            rec["reduce_input_size"] = True
            rec["suggested_input_size"] = (640, 640)
            self.degradation_level += 1
            logger.error("Memory reduction: EXTREME pressure - input size 640x640, batch size 1")
        return rec

    def handle_oom_error(self) -> Dict[str, Any]:
        self.oom_events += 1
        logger.error("GPU OOM detected (event #%d) - attempting recovery", self.oom_events)
        self.clear_cache()
        self.degradation_level = min(self.degradation_level + 1, 5)
        return {"clear_cache": True, "reduce_batch_size": True, "reduce_input_size": True,
                "cpu_fallback": False,  # no CPU path in the B200 build, however many OOM events pile up
                "suggested_input_size": (640, 640), "suggested_batch_size": 1, "wait_time": 1.0}

    def record_recovery(self):
        self.recoveries += 1

    def get_memory_stats(self) -> Dict[str, Any]:
        stats = {"cuda_available": self.cuda_available, "current_pressure": self.current_pressure.value,
                 "oom_events": self.oom_events, "recoveries": self.recoveries,
                 "degradation_level": self.degradation_level}
        if self.cuda_available:
            try:
                t = _cuda()
                used, total = self._usage()
                alloc = t.cuda.memory_allocated(self.device)
                stats.update({"allocated_mb": alloc / 1024 / 1024, "allocated_gb": alloc / 1e9,
                              "reserved_mb": used / 1024 / 1024, "reserved_gb": used / 1e9,
                              "total_mb": total / 1024 / 1024, "total_gb": total / 1e9,
                              "usage_percent": used / total * 100 if total else 0.0,
                              "allocated_percent": alloc / total * 100 if total else 0.0})
            except Exception as e:
                logger.error("Failed to get GPU memory stats: %s", e)
        return stats

    def clear_cache(self):
        t = _cuda()
        if t is not None:
            try:
                t.cuda.empty_cache()  # releases only what torch itself cached; libtcd arenas are static
            except Exception as e:
                logger.error("Failed to clear GPU cache: %s", e)

    def reset_degradation(self):
        self.degradation_level = 0
