"""BASELINE.json configs[3] -- kept as a thin entry point: the measurement lives in bench.py (`--workload prefill`).

  python tools/prefill_bench.py [--layers 30] [--no-fused-epilogue]
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = ["--workload", "prefill"]
if "--layers" in sys.argv:
    args += ["--prefill-layers", sys.argv[sys.argv.index("--layers") + 1]]
if "--no-fused-epilogue" in sys.argv:
    args += ["--no-fused-epilogue"]
sys.exit(subprocess.call([sys.executable, os.path.join(ROOT, "bench.py")] + args))
