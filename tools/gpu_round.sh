#!/bin/bash
# one gpurun call: tests, smoke, bench, FP64 peak, QED step, then the per-kernel ncu captures (tag = $1)
T=${1:-r02}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${T}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"
tools/fp64_peak > $O/${T}_fp64.json 2>&1
python tools/bench_qed.py --out $O/${T}_qed.json > $O/${T}_qed.log 2>&1; echo "qed rc=$?"
if [ "$2" != "noncu" ]; then
NCU="ncu --set full --clock-control none --import-source on"
python tools/profile_kernels.py --phase lean_e --steps 6 > $O/${T}_prof_plain.log 2>&1 &&
$NCU -k regex:k_push_lean_fast -s 4 -c 1 -f -o $O/${T}_push_lean_e python tools/profile_kernels.py --phase lean_e --steps 6 > $O/${T}_ncu1.log 2>&1
python tools/profile_kernels.py --phase lean_i --steps 6 >> $O/${T}_prof_plain.log 2>&1 &&
$NCU -k regex:k_push_lean_fast -s 4 -c 1 -f -o $O/${T}_push_lean_i python tools/profile_kernels.py --phase lean_i --steps 6 > $O/${T}_ncu3.log 2>&1
python tools/profile_kernels.py --phase qed >> $O/${T}_prof_plain.log 2>&1 &&
$NCU -k 'regex:^(k_push_full_fast|k_scatter|k_photon_push|k_emit)$' -s 4 -c 4 -f -o $O/${T}_qed python tools/profile_kernels.py --phase qed > $O/${T}_ncu4.log 2>&1
python tools/profile_kernels.py --phase two_stream >> $O/${T}_prof_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_two_stream_launches.csv python tools/profile_kernels.py --phase two_stream > $O/${T}_ncu5.log 2>&1
python bench.py --steps 3 --warmup 2 --reps 1 --no-cpu-baseline --no-parity --no-config5 > $O/${T}_bench_small.json 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/${T}_bench_launches.csv python bench.py --steps 3 --warmup 2 --reps 1 --no-cpu-baseline --no-parity --no-config5 > $O/${T}_ncu6.log 2>&1
python tools/bench_qed.py --cpu-nx 0 --steps 4 --warmup 2 --out $O/${T}_qed_small.json > /dev/null 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/${T}_qed_launches.csv python tools/bench_qed.py --cpu-nx 0 --steps 4 --warmup 2 --out $O/${T}_qed_small.json > $O/${T}_ncu7.log 2>&1
fi
ls -la $O | tail -30
tail -3 $O/${T}_pytest.log; cat $O/${T}_smoke.log | tail -2; cat $O/${T}_bench.json; cat $O/${T}_fp64.json; tail -2 $O/${T}_qed.log
