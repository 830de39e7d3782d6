for m in rows stream mma twopass; do echo "== $m"; MINIPS_DENSE_LR=$m timeout 300 python tools/profile_dense_lr.py 2>&1 | tail -1; done
