#!/usr/bin/env bash
# Same calling convention as the reference's hwN/run.sh (hw1..hw5: 2 args; hw6..hw8: 5 or 6 args,
# hw8/run.sh:2-9 including its two echo lines).
HERE="$(cd "$(dirname "$0")" && pwd)"
BIN="$HERE/raytracing_course_hw_b200/lib/b200rt"
if [ $# -eq 6 ]; then
    echo "Launching version with environment map"
    "$BIN" "$1" "$2" "$3" "$4" "$5" "$6"
elif [ $# -eq 5 ]; then
    echo "Launching version without environment map"
    "$BIN" "$1" "$2" "$3" "$4" "$5"
else
    "$BIN" "$@"
fi
