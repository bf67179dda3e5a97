"""Run every GPU test *function* in its own process with a timeout, keep going on failure.

A device-side trap poisons the CUDA context of its process; isolating functions means one
broken kernel does not hide the state of the others.  Writes a summary to stdout.
usage: python scripts/run_isolated.py tests/test_gpu_kernels.py [more files] [--timeout 300]
"""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["PYTHONPATH"] = ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")

args = [a for a in sys.argv[1:] if not a.startswith("--")]
timeout = 300
if "--timeout" in sys.argv:
    timeout = int(sys.argv[sys.argv.index("--timeout") + 1])
    args = [a for a in args if a != str(timeout)]

col = subprocess.run([sys.executable, "-m", "pytest", "--collect-only", "-q", "-m", "gpu", *args],
                     capture_output=True, text=True)
funcs = []
for line in col.stdout.splitlines():
    if "::" in line:
        f = line.split("[")[0].strip()
        if f not in funcs:
            funcs.append(f)
print(f"collected {len(funcs)} test functions", flush=True)
summary = []
for f in funcs:
    t0 = time.time()
    try:
        r = subprocess.run([sys.executable, "-m", "pytest", f, "-q", "-m", "gpu", "-x", "--no-header", "-p", "no:cacheprovider"],
                           capture_output=True, text=True, timeout=timeout)
        ok = r.returncode == 0
        tail = "\n".join((r.stdout + r.stderr).splitlines()[-25:])
    except subprocess.TimeoutExpired as e:
        ok, tail = False, f"TIMEOUT after {timeout}s\n" + str((e.stdout or b"")[-2000:])
    dt = time.time() - t0
    summary.append((f, ok, dt))
    last = [l for l in (r.stdout if 'r' in dir() else '').splitlines() if l.strip()][-1:] if ok else []
    print(f"=== {'PASS' if ok else 'FAIL'} {f} ({dt:.1f}s) {' '.join(last)}", flush=True)
    if not ok:
        print(tail, flush=True)
print("\nSUMMARY")
for f, ok, dt in summary:
    print(f"{'PASS' if ok else 'FAIL'} {dt:6.1f}s {f}")
sys.exit(0 if all(ok for _, ok, _ in summary) else 1)
