"""Where does pinned host memory land relative to the GPU?  H2D bandwidth of a 1 GB pinned buffer allocated and first
touched with the process bound to each NUMA node's CPUs."""
import glob, os, time
import torch
def cpulist(s):
    out = []
    for part in s.strip().split(","):
        if not part: continue
        a, _, b = part.partition("-")
        out += list(range(int(a), int(b or a) + 1))
    return out
allowed = sorted(os.sched_getaffinity(0))
print("allowed cpus", allowed)
bus = torch.cuda.get_device_properties(0)
pci = None
try:
    import subprocess
    pci = subprocess.check_output(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", "0"], text=True).strip().lower()
    if pci.startswith("00000000:"): pci = pci[4:]
    print("gpu pci", pci, "numa_node", open(f"/sys/bus/pci/devices/{pci}/numa_node").read().strip())
except Exception as e:
    print("pci lookup failed", e)
nodes = {}
for d in sorted(glob.glob("/sys/devices/system/node/node[0-9]*")):
    n = int(d.rsplit("node", 1)[1])
    nodes[n] = [c for c in cpulist(open(d + "/cpulist").read()) if c in allowed]
    print("node", n, "allowed cpus", nodes[n][:8], "..." if len(nodes[n]) > 8 else "")
dev = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
for n, cpus in nodes.items():
    if not cpus: continue
    os.sched_setaffinity(0, cpus)
    host = torch.empty(1 << 30, dtype=torch.uint8).pin_memory()
    host.fill_(1)
    torch.cuda.synchronize()
    for rep in range(2):
        t0 = time.perf_counter(); dev.copy_(host, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"node {n}: H2D {1.0737 / dt:.1f} GB/s")
    del host
os.sched_setaffinity(0, allowed)
