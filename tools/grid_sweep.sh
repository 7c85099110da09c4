run() { echo -n "$1 : "; for d in 60 110; do CABM_GRID_MULT=$1 python tools/profile_day.py --start-day $d --days 20 --graph | python -c "import sys,json; print(json.load(sys.stdin)['device_us_per_day'], end=' ')"; done; echo; }
run 8,8,8,4,8,1,2,2,4
run 4,8,8,4,8,1,2,2,4
run 2,8,8,4,8,1,2,2,4
run 1,8,8,4,8,1,2,2,4
run 8,4,8,4,8,1,2,2,4
run 8,2,8,4,8,1,2,2,4
run 8,8,4,4,8,1,2,2,4
run 8,8,8,2,8,1,2,2,4
run 8,8,8,1,8,1,2,2,4
run 8,8,8,4,4,1,2,2,4
run 8,8,8,4,2,1,2,2,4
run 8,8,8,4,8,2,2,2,4
run 8,8,8,4,8,1,2,1,4
run 8,8,8,4,8,1,2,2,2
run 8,8,8,4,8,1,2,2,1
