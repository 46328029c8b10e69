#!/bin/bash
# Round measurement set, run on the GPU box:  gpurun --timeout 2400 -- 'bash tools/measure_round.sh r01'
# 1. bench.py both arms (never under a profiler)  2. ncu launch list of the same bench command
# 3. one full ncu capture each of k_wave (both sweeps), k_cost_fast, k_bm_fast  4. small-section capture of one whole step
# Everything lands in gpurun_out/ (scratch); tools/make_profiles.py turns it into profiles/.
# gpurun copies back at most 64 MiB: run it in parts (second argument: bench | ncu | all; default everything)
TAG=${1:-r02}
PART=${2:-everything}
O=gpurun_out
mkdir -p $O
want() { [ "$PART" = everything ] || [ "$PART" = "$1" ]; }
if want bench; then
python bench.py > $O/bench_$TAG.json 2> $O/bench_$TAG.err
python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref_$TAG.json 2>> $O/bench_$TAG.err
python tools/bench_configs.py > $O/configs_$TAG.jsonl 2>> $O/bench_$TAG.err
fi
NCU="ncu --clock-control none"
if want bench; then
$NCU --metrics gpu__time_duration.sum -c 600 --csv --log-file $O/launches_$TAG.csv \
    python bench.py --steps 2 --warmup 3 --batch 128 --no-cpu --no-parity > $O/ncu.log 2>&1
fi
if want ncu; then
$NCU --set full --import-source on -k regex:k_wave8 -s 2 -c 2 -o $O/wave_$TAG -f python tools/run_sgbm_once.py 128 2 > $O/ncu_wave.log 2>&1
$NCU --set full --import-source on -k regex:k_cost -s 1 -c 1 -o $O/cost_$TAG -f python tools/run_sgbm_once.py 128 2 > $O/ncu_cost.log 2>&1
$NCU --set full --import-source on -k regex:k_bm_fast -s 1 -c 1 -o $O/bm_$TAG -f python tools/bench_configs.py cfg2 > $O/ncu_bm.log 2>&1
fi
# (No compute-sanitizer pass: the GPU pool refuses it -- "compute-sanitizer is closed on this pool and stays closed",
#  exit code 86, recorded in profiles/README.md.  Races and bad accesses in the band pipeline are hunted with the parity
#  suites instead: small shapes around every band / chunk edge, batch == singles, repeated calls, all against cv2.)
if want all; then
$NCU --section SpeedOfLight --section LaunchStats --section Occupancy --section MemoryWorkloadAnalysis -s 30 -c 30 -o $O/all_$TAG -f \
    python bench.py --steps 1 --warmup 3 --batch 32 --no-cpu --no-parity > $O/ncu_all.log 2>&1
fi
du -sh $O; cut -c1-600 $O/bench_$TAG.json
