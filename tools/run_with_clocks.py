#!/usr/bin/env python
"""Runs a measurement command while sampling the GPU's SM clock and throttle reasons (the same nvidia-smi poll bench.py
uses) and prints ONE JSON document: the command's own JSON output plus a "clocks" record.

    python tools/run_with_clocks.py tools/micro/red_probe > profiles/r02_red_microbench.json
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import ClockSampler  # noqa: E402


def main():
    cmd = sys.argv[1:]
    if not cmd:
        raise SystemExit("usage: run_with_clocks.py <command> [args...]")
    sampler = ClockSampler(0)
    sampler.start()
    time.sleep(0.5)
    sampler.mark()
    outs = []
    t0 = time.time()
    while time.time() - t0 < 1.5 or not outs:  # repeat short probes so that the 200 ms poll sees them under load
        outs.append(subprocess.run(cmd, stdout=subprocess.PIPE, text=True, check=True).stdout)
    clocks = sampler.stop()
    try:
        doc = json.loads(outs[-1])
    except ValueError:
        doc = {"output": outs[-1]}
    doc["clocks"] = clocks
    doc["repeats"] = len(outs)
    print(json.dumps(doc, indent=1))


if __name__ == "__main__":
    main()
