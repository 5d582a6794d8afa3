#!/bin/bash
# One gpurun call: GPU parity tests, smoke, a bench line, the ncu launch list and one --set full capture
# of the kernels named in $1 (regex, default: the FAST kernel).  Outputs land in gpurun_out/.
set -u
K=${1:-fast_score_kernel}
TAG=${2:-r1}
SKIP=${3:-6}
CNT=${4:-8}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; cat gpurun_out/${TAG}_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err; cat gpurun_out/${TAG}_bench_ref.json
P="python bench.py --steps 2 --warmup 1 --pairs-per-step 64 --pool 64 --skip-e2e --no-cpu-baseline"
$P > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${TAG}_launches.csv $P > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu list rc=$?"
$P > gpurun_out/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$K" -s $SKIP -c $CNT -f -o gpurun_out/${TAG}_prof $P > gpurun_out/${TAG}_ncu2.log 2>&1
echo "ncu full rc=$?"
