"""Host-side placement for the host-buffer entry points (exo_step_host): pin the calling process to the CPU cores / NUMA
node nearest its GPU BEFORE it allocates pinned buffers, so the 21 MB of observations each step crosses one PCIe root
complex into local DRAM instead of the socket interconnect. With 8 ranks on one box the default placement (every rank on
NUMA node 0) made the aggregate device->host rate, not the kernels, the limit (SCALE_r01: e2e efficiency 0.28 at N = 8)."""
import os
import re
import subprocess


def _parse_cpu_list(text):
    cpus = set()
    for part in text.split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        else:
            cpus.add(int(part))
    return cpus


def gpu_affinity(index):
    """(cpu set, numa string) of GPU `index` from `nvidia-smi topo -m` ("CPU Affinity" / "NUMA Affinity" columns);
    (None, None) if it cannot be read."""
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
    except (OSError, subprocess.SubprocessError):
        return None, None
    out = re.sub(r"\x1b\[[0-9;]*m", "", out)
    header = None
    for line in out.splitlines():
        cols = [c for c in line.split("\t") if c != ""]
        if not cols:
            continue
        if header is None and any("CPU Affinity" in c for c in cols):
            header = [c.strip() for c in cols]
            continue
        if header is not None and cols[0].strip() == "GPU%d" % index:
            vals = [c.strip() for c in cols]
            names = [""] + header if len(header) + 1 == len(vals) else header      # the first column has no header
            rec = dict(zip(names, vals))
            cpu = rec.get("CPU Affinity")
            try:
                return (_parse_cpu_list(cpu) if cpu else None), rec.get("NUMA Affinity")
            except ValueError:
                return None, rec.get("NUMA Affinity")
    return None, None


def pin_to_gpu(index):
    """sched_setaffinity to the GPU's CPU affinity set (intersected with what the process may use). Returns a small dict
    for logging; never raises."""
    info = {"gpu": index, "pinned": False}
    try:
        allowed = os.sched_getaffinity(0)
        cpus, numa = gpu_affinity(index)
        info["numa"] = numa
        if cpus:
            use = cpus & allowed
            if use and use != allowed:
                os.sched_setaffinity(0, use)
                info["pinned"] = True
            info["cpus"] = len(use or allowed)
        else:
            info["cpus"] = len(allowed)
    except (AttributeError, OSError) as ex:
        info["error"] = str(ex)
    return info
