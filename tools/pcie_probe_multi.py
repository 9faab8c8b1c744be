"""Host<->device ceiling with N GPUs busy at once: one process per GPU, all copying pinned buffers for
the same wall-clock window.  Prints per-GPU and aggregate GB/s for H2D alone, D2H alone and both directions together, plus
the topology the numbers have to be read against.

    python tools/pcie_probe_multi.py 8            # parent: spawns one worker per GPU
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WINDOW_S = 2.0
CHUNK = 256 << 20


def worker(gpu: int, t_start: float):
    import torch

    torch.cuda.set_device(gpu)
    h_in = torch.empty(CHUNK, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(CHUNK, dtype=torch.uint8).pin_memory()
    d_a = torch.empty(CHUNK, dtype=torch.uint8, device="cuda")
    d_b = torch.empty(CHUNK, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    res = {"gpu": gpu}
    for mode in ("h2d", "d2h", "both"):
        t_start += WINDOW_S + 1.0
        for _ in range(2):  # warm-up
            d_a.copy_(h_in, non_blocking=True)
            h_out.copy_(d_b, non_blocking=True)
        torch.cuda.synchronize()
        while time.time() < t_start:
            pass
        n1 = n2 = 0
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < WINDOW_S:
            if mode in ("h2d", "both"):
                with torch.cuda.stream(s1):
                    d_a.copy_(h_in, non_blocking=True)
                n1 += 1
            if mode in ("d2h", "both"):
                with torch.cuda.stream(s2):
                    h_out.copy_(d_b, non_blocking=True)
                n2 += 1
            s1.synchronize()
            s2.synchronize()
        dt = time.perf_counter() - t0
        res[mode] = {"h2d_gbs": n1 * CHUNK / dt / 1e9, "d2h_gbs": n2 * CHUNK / dt / 1e9}
    print("RESULT " + json.dumps(res), flush=True)


def main():
    if len(sys.argv) >= 3 and sys.argv[1] == "--worker":
        worker(int(sys.argv[2]), float(sys.argv[3]))
        return
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    t_start = time.time() + 25.0  # time for every worker to import torch and pin its buffers
    procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--worker", str(g), str(t_start)], stdout=subprocess.PIPE,
                              text=True) for g in range(n)]
    outs = [p.communicate()[0] for p in procs]
    rows = [json.loads(l[7:]) for o in outs for l in o.splitlines() if l.startswith("RESULT ")]
    rows.sort(key=lambda r: r["gpu"])
    summary = {"n_gpus": n, "per_gpu": rows, "aggregate": {}}
    for mode in ("h2d", "d2h", "both"):
        summary["aggregate"][mode] = {"h2d_gbs": sum(r[mode]["h2d_gbs"] for r in rows), "d2h_gbs": sum(r[mode]["d2h_gbs"] for r in rows)}
    try:
        summary["topo"] = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=30).stdout
        summary["numa_nodes"] = sorted(d for d in os.listdir("/sys/devices/system/node") if d.startswith("node"))
        summary["cpus"] = os.cpu_count()
    except Exception as e:
        summary["topo_error"] = str(e)
    print(json.dumps(summary, indent=1))


if __name__ == "__main__":
    main()
