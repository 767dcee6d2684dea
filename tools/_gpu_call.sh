set -x
cd $GRAFT_REPO_ROOT
timeout 120 python tools/probe_run.py 4 2>&1 | tail -6
