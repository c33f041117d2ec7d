#!/usr/bin/env python3
"""stdin: output of tools/profile_step.py --counters; prints the counters and stage times on one line."""
import ast, sys
for line in sys.stdin:
    if line.startswith("{'kernel_ms'"):
        d = ast.literal_eval(line)
        print(f"ms {d['kernel_ms']:.2f} ext {d['extend_ms']:.2f} shade {d['shade_ms']:.2f} light {d['light_ms']:.2f} | "
              f"node {d['node_tests'] / 1e6:.0f}M tri {d['tri_tests'] / 1e6:.0f}M lnode {d['light_node_tests'] / 1e6:.0f}M ltri {d['light_tri_tests'] / 1e6:.0f}M")
