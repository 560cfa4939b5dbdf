#!/bin/bash
echo "== default"; python tools/game_frame_run.py 64 8 6 | grep -E "build_lines"
for v in openspades_b200/lib/var/libosb200_*.so; do echo "== $v"; OSB200_LIB=$PWD/$v python tools/game_frame_run.py 64 8 6 | grep -E "build_lines"; done
