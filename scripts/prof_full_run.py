"""Runs ONE timed window (2 steps) of each named bench configuration in a single process, so that one
`ncu --set full --profile-from-start off` invocation sees every hot-path kernel of the BASELINE configs
(scripts/prof_full.sh).  Numbers printed under ncu are not bench values."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench

for cfg in sys.argv[1:] or ["orr"]:
    sys.argv = ["bench.py", "--config", cfg, "--steps", "2", "--warmup", "3", "--no-e2e", "--no-baseline",
                "--min-timed-ms", "0.001"]
    if cfg == "orr":
        sys.argv += ["--hyper", "solveEvery=2"]      # the weight refresh falls inside the 2-step window
    bench.main()
