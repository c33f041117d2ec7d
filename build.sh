#!/usr/bin/env bash
# Counterpart of the reference's hwN/build.sh: builds the CUDA core (sm_100a), the host library and the CLI.
HERE="$(cd "$(dirname "$0")" && pwd)"
cd "$HERE" && python -c "from raytracing_course_hw_b200 import build; build.build_cuda(); build.build_host(); build.build_cli()"
