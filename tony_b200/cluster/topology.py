"""GPU / NVLink topology and GPU metrics discovery without importing torch.

Two consumers: the local resource manager (which GPUs exist, which CPUs sit next to each)
and the TaskMonitor (utilisation + memory gauges).  The reference forks ``nvidia-smi -x -q``
and parses the XML with JAXB (tony-core/src/main/java/com/linkedin/tony/util/gpu/
GpuDiscoverer.java:52-56,88-190, GpuDeviceInformation*.java, PerGpu*.java); the same XML is
parsed here with ElementTree (fixture: the reference's sample is the shape we accept), plus
``nvidia-smi topo -m`` / sysfs for the NVLink matrix and CPU affinity that a one-task-per-GPU
placement on an NVSwitch box needs.
"""
from __future__ import annotations

import json
import os
import re
import shutil
import subprocess
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import Dict, List, Optional

DEFAULT_BINARY = "nvidia-smi"
SEARCH_DIRS = ("/usr/bin", "/bin", "/usr/local/nvidia/bin", "/usr/local/cuda/bin")
EXEC_TIMEOUT_S = 10
MAX_REPEATED_ERRORS = 10


@dataclass
class PerGpuInfo:
    product_name: str = ""
    uuid: str = ""
    minor_number: int = -1
    gpu_util_pct: float = 0.0
    fb_total_mib: float = 0.0
    fb_used_mib: float = 0.0
    bar1_total_mib: float = 0.0
    bar1_used_mib: float = 0.0
    temperature_c: float = 0.0

    @property
    def fb_usage_pct(self) -> float:
        return 100.0 * self.fb_used_mib / self.fb_total_mib if self.fb_total_mib else 0.0

    @property
    def bar1_usage_pct(self) -> float:
        return 100.0 * self.bar1_used_mib / self.bar1_total_mib if self.bar1_total_mib else 0.0


@dataclass
class GpuDeviceInformation:
    driver_version: str = ""
    gpus: List[PerGpuInfo] = field(default_factory=list)


def _num(text: Optional[str]) -> float:
    if not text:
        return 0.0
    m = re.search(r"-?\d+(\.\d+)?", text)
    return float(m.group(0)) if m else 0.0


def parse_smi_xml(xml_text: str) -> GpuDeviceInformation:
    """Parse ``nvidia-smi -x -q`` output."""
    root = ET.fromstring(xml_text)
    info = GpuDeviceInformation(driver_version=(root.findtext("driver_version") or "").strip())
    for g in root.findall("gpu"):
        p = PerGpuInfo()
        p.product_name = (g.findtext("product_name") or "").strip()
        p.uuid = (g.findtext("uuid") or "").strip()
        p.minor_number = int(_num(g.findtext("minor_number"))) if g.findtext("minor_number") else -1
        p.gpu_util_pct = _num(g.findtext("utilization/gpu_util"))
        p.fb_total_mib = _num(g.findtext("fb_memory_usage/total"))
        p.fb_used_mib = _num(g.findtext("fb_memory_usage/used"))
        p.bar1_total_mib = _num(g.findtext("bar1_memory_usage/total"))
        p.bar1_used_mib = _num(g.findtext("bar1_memory_usage/used"))
        p.temperature_c = _num(g.findtext("temperature/gpu_temp"))
        info.gpus.append(p)
    return info


class NvmlSampler:
    """GPU gauges straight from ``libnvidia-ml.so.1`` through ctypes: a few microseconds per sample
    instead of forking ``nvidia-smi -x -q`` (about 100 ms, per task, per interval).  Only the calls the
    monitor needs; any failure makes ``GpuDiscoverer`` fall back to the XML path."""

    class _Util(__import__("ctypes").Structure):
        _fields_ = [("gpu", __import__("ctypes").c_uint), ("memory", __import__("ctypes").c_uint)]

    class _Mem(__import__("ctypes").Structure):
        _fields_ = [("total", __import__("ctypes").c_ulonglong), ("free", __import__("ctypes").c_ulonglong),
                    ("used", __import__("ctypes").c_ulonglong)]

    class _Bar1(__import__("ctypes").Structure):
        _fields_ = [("bar1Total", __import__("ctypes").c_ulonglong), ("bar1Free", __import__("ctypes").c_ulonglong),
                    ("bar1Used", __import__("ctypes").c_ulonglong)]

    def __init__(self):
        import ctypes

        self.ct = ctypes
        self.lib = ctypes.CDLL("libnvidia-ml.so.1")
        if self.lib.nvmlInit_v2() != 0:
            raise OSError("nvmlInit_v2 failed")
        n = ctypes.c_uint()
        if self.lib.nvmlDeviceGetCount_v2(ctypes.byref(n)) != 0:
            raise OSError("nvmlDeviceGetCount_v2 failed")
        self.handles = []
        for i in range(n.value):
            h = ctypes.c_void_p()
            if self.lib.nvmlDeviceGetHandleByIndex_v2(ctypes.c_uint(i), ctypes.byref(h)) != 0:
                raise OSError("nvmlDeviceGetHandleByIndex_v2 failed")
            self.handles.append(h)

    def sample(self) -> GpuDeviceInformation:
        ct = self.ct
        info = GpuDeviceInformation()
        for i, h in enumerate(self.handles):
            p = PerGpuInfo(minor_number=i)
            u, m, b, t = self._Util(), self._Mem(), self._Bar1(), ct.c_uint()
            if self.lib.nvmlDeviceGetUtilizationRates(h, ct.byref(u)) == 0:
                p.gpu_util_pct = float(u.gpu)
            if self.lib.nvmlDeviceGetMemoryInfo(h, ct.byref(m)) == 0:
                p.fb_total_mib, p.fb_used_mib = m.total / 2 ** 20, m.used / 2 ** 20
            if self.lib.nvmlDeviceGetBAR1MemoryInfo(h, ct.byref(b)) == 0:
                p.bar1_total_mib, p.bar1_used_mib = b.bar1Total / 2 ** 20, b.bar1Used / 2 ** 20
            if self.lib.nvmlDeviceGetTemperature(h, ct.c_int(0), ct.byref(t)) == 0:
                p.temperature_c = float(t.value)
            mn = ct.c_uint()
            if self.lib.nvmlDeviceGetMinorNumber(h, ct.byref(mn)) == 0:
                p.minor_number = int(mn.value)
            info.gpus.append(p)
        return info


class GpuDiscoverer:
    """GPU gauges for the TaskMonitor: NVML in-process when the library loads (TONY_B200_NVML=0 disables),
    else the reference's way - find the nvidia-smi binary once, fork ``-x -q`` and parse the XML; gives
    up after repeated errors."""

    def __init__(self, exec_path: Optional[str] = None):
        self.binary = self._locate(exec_path)
        self.errors = 0
        self.nvml: Optional[NvmlSampler] = None
        if exec_path is None and os.environ.get("TONY_B200_NVML", "1") != "0" \
                and os.environ.get("TONY_B200_FAKE_GPUS") is None:
            try:
                self.nvml = NvmlSampler()
            except (OSError, AttributeError):
                self.nvml = None

    @staticmethod
    def _locate(exec_path: Optional[str]) -> Optional[str]:
        if exec_path:
            cand = exec_path if os.path.isfile(exec_path) else os.path.join(exec_path, DEFAULT_BINARY)
            return cand if os.path.isfile(cand) else None
        found = shutil.which(DEFAULT_BINARY)
        if found:
            return found
        for d in SEARCH_DIRS:
            cand = os.path.join(d, DEFAULT_BINARY)
            if os.path.isfile(cand):
                return cand
        return None

    @property
    def usable(self) -> bool:
        return (self.nvml is not None or self.binary is not None) and self.errors < MAX_REPEATED_ERRORS

    def sample(self) -> Optional[GpuDeviceInformation]:
        if not self.usable:
            return None
        if self.nvml is not None:
            try:
                return self.nvml.sample()
            except Exception:  # noqa: BLE001 - fall back to the XML path for good
                self.nvml = None
        if self.binary is None:
            return None
        try:
            out = subprocess.run([self.binary, "-x", "-q"], capture_output=True, text=True,
                                 timeout=EXEC_TIMEOUT_S, check=True).stdout
            info = parse_smi_xml(out)
            self.errors = 0
            return info
        except (OSError, subprocess.SubprocessError, ET.ParseError):
            self.errors += 1
            return None


def discover_gpu_indices() -> List[int]:
    """Indices of the GPUs on this node ([] on a CPU-only box)."""
    env = os.environ.get("TONY_B200_FAKE_GPUS")
    if env is not None:
        return list(range(int(env)))
    binary = GpuDiscoverer._locate(None)
    if not binary:
        return []
    try:
        out = subprocess.run([binary, "--query-gpu=index", "--format=csv,noheader"], capture_output=True,
                             text=True, timeout=EXEC_TIMEOUT_S, check=True).stdout
        return [int(x) for x in out.split() if x.strip().isdigit()]
    except (OSError, subprocess.SubprocessError, ValueError):
        return []


def physical_memory_mb() -> int:
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemTotal:"):
                    return int(line.split()[1]) // 1024
    except OSError:
        pass
    return 64 * 1024


def _gpu_bus_ids() -> Dict[int, str]:
    binary = GpuDiscoverer._locate(None)
    if not binary:
        return {}
    try:
        out = subprocess.run([binary, "--query-gpu=index,pci.bus_id", "--format=csv,noheader"],
                             capture_output=True, text=True, timeout=EXEC_TIMEOUT_S, check=True).stdout
    except (OSError, subprocess.SubprocessError):
        return {}
    ids = {}
    for line in out.splitlines():
        parts = [p.strip() for p in line.split(",")]
        if len(parts) == 2 and parts[0].isdigit():
            ids[int(parts[0])] = parts[1]
    return ids


def cpus_near_gpus(gpu_ids: List[int]) -> Optional[str]:
    """CPU list (taskset syntax) local to the given GPUs, from sysfs ``local_cpulist``."""
    bus = _gpu_bus_ids()
    lists = []
    for g in gpu_ids:
        b = bus.get(g)
        if not b:
            return None
        # nvidia-smi prints an 8-digit domain, sysfs uses 4
        b = b.lower()
        if len(b.split(":")[0]) == 8:
            b = b[4:]
        try:
            with open(f"/sys/bus/pci/devices/{b}/local_cpulist") as f:
                lists.append(f.read().strip())
        except OSError:
            return None
    lists = [l for l in lists if l]
    return ",".join(dict.fromkeys(lists)) if lists else None


def nvlink_topology() -> Dict[str, object]:
    """NVLink / NVSwitch connectivity as JSON-able data (exported to tasks as TONY_TOPOLOGY)."""
    topo: Dict[str, object] = {"gpus": discover_gpu_indices(), "matrix": [], "nvswitch": False}
    binary = GpuDiscoverer._locate(None)
    if not binary or not topo["gpus"]:
        return topo
    try:
        out = subprocess.run([binary, "topo", "-m"], capture_output=True, text=True,
                             timeout=EXEC_TIMEOUT_S, check=True).stdout
    except (OSError, subprocess.SubprocessError):
        return topo
    rows = []
    for line in out.splitlines():
        cells = line.split()
        if cells and re.fullmatch(r"GPU\d+", cells[0]):
            rows.append([c for c in cells[1:] if re.fullmatch(r"X|NV\d+|SYS|NODE|PHB|PXB|PIX", c)])
    n = len(rows)
    topo["matrix"] = [r[:n] for r in rows]
    # every pair connected with the same NVn count == switched fabric (NV18 on B200 HGX)
    links = {c for r in topo["matrix"] for c in r if c.startswith("NV")}
    topo["nvswitch"] = n > 1 and len(links) == 1 and all(
        c == "X" or c.startswith("NV") for r in topo["matrix"] for c in r)
    topo["links_per_pair"] = next(iter(links), None)
    return topo


def topology_json() -> str:
    return json.dumps(nvlink_topology(), separators=(",", ":"))
