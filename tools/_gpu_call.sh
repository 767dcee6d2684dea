set -x
cd $GRAFT_REPO_ROOT
timeout 80 python tools/vs_one_rank.py --rank 1 --slices 0 --frames 120 2>&1 | grep "^frame" | sed -n '1p;3p;60p;120p' | cut -c1-140
