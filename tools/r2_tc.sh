timeout 180 python tools/tc_check.py 2>&1 | tail -30
echo "rc=$?"
