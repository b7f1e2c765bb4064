#!/bin/bash
# Copies the reference's OWN Python package (unmodified) to oracle/_ref/ so that it can travel to the GPU box with the
# gpurun snapshot: bench.py --impl reference then times the unmodified bluesky_gym.envs.HorizontalCREnv class (on top of
# oracle/refshim's BlueSky stand-in) instead of the port in oracle/envs_np.py.  oracle/_ref/ is git-ignored: reference
# sources never enter the history.  Run HERE (needs /root/reference); __graft_entry__.build() runs it when it can.
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
SRC="${1:-/root/reference}"
[ -d "$SRC/bluesky_gym" ] || { echo "make_oracle_ref: $SRC/bluesky_gym not found (nothing copied)"; exit 0; }
rm -rf "$ROOT/oracle/_ref"
mkdir -p "$ROOT/oracle/_ref"
cp -r "$SRC/bluesky_gym" "$ROOT/oracle/_ref/bluesky_gym"
find "$ROOT/oracle/_ref" -name "__pycache__" -type d -prune -exec rm -rf {} +
( cd "$SRC" && find bluesky_gym -name "*.py" -type f | sort | xargs sha256sum ) > "$ROOT/oracle/_ref/SHA256SUMS"
echo "make_oracle_ref: $(find "$ROOT/oracle/_ref/bluesky_gym" -name '*.py' | wc -l) files copied to oracle/_ref/bluesky_gym"
