#!/bin/bash
# Round-2 evidence run on ONE B200 (tools/gpu.sh 1500 'bash tools/r2_profile.sh'): probes first (plain runs), then the
# ncu passes -- each profiled command has just exited 0 without ncu.  Reports land in gpurun_out/ (scratch);
# tools/ncu_summary.py turns them into the text summaries committed under profiles/.
set -u
mkdir -p gpurun_out
O=gpurun_out
ROOT=$(pwd)
python tools/band_probe.py 8 > $O/r2o_band_probe.jsonl 2> $O/r2o_band_probe.err; echo "band rc=$?"
python tools/dyn_probe.py > $O/r2o_dyn_probe.jsonl 2> $O/r2o_dyn_probe.err; echo "dyn rc=$?"
NCU="ncu --clock-control none"
# launch list of the headline command, wavefront kernels pinned (under ncu the timed choice sees serialised launches)
python bench.py --kernel 1 --steps 1 --warmup 3 --no-cpu-baseline --no-c4-tiles > $O/r2o_bench_k1.json 2> $O/r2o_bench_k1.err; echo "bench k1 rc=$?"
$NCU --metrics gpu__time_duration.sum -s 12000 -c 400 --csv --log-file $O/r2o_launches_C3.csv \
    python bench.py --kernel 1 --steps 1 --warmup 3 --no-cpu-baseline --no-c4-tiles > $O/r2o_ncu_launches.log 2>&1; echo "launch list rc=$?"
# full captures: one mid-frame launch of each wavefront kernel (4-spp frame, one row group), the pooled and the streaming kernel
python tools/prof_one.py C3 4 kernel=1 wf_groups=1 > $O/r2o_prof_wf.log 2>&1; echo "prof wf plain rc=$?"
$NCU --set full --import-source on -k regex:wf_traverse -s 45 -c 1 -f -o $O/r2o_wf_traverse python tools/prof_one.py C3 4 kernel=1 wf_groups=1 > $O/r2o_ncu_wf_traverse.log 2>&1; echo "ncu traverse rc=$?"
$NCU --set full --import-source on -k regex:wf_shade -s 45 -c 1 -f -o $O/r2o_wf_shade python tools/prof_one.py C3 4 kernel=1 wf_groups=1 > $O/r2o_ncu_wf_shade.log 2>&1; echo "ncu shade rc=$?"
python tools/prof_one.py C3 2 kernel=3 > $O/r2o_prof_pool.log 2>&1; echo "prof pool plain rc=$?"
timeout 600 $NCU --set full --import-source on -k regex:pool_render -s 1 -c 1 -f -o $O/r2o_pool python tools/prof_one.py C3 2 kernel=3 > $O/r2o_ncu_pool.log 2>&1; echo "ncu pool rc=$?"
python tools/prof_one.py C3 2 kernel=4 > $O/r2o_prof_stream.log 2>&1; echo "prof stream plain rc=$?"
timeout 600 $NCU --set full --import-source on -k regex:stream_render -s 1 -c 1 -f -o $O/r2o_stream python tools/prof_one.py C3 2 kernel=4 > $O/r2o_ncu_stream.log 2>&1; echo "ncu stream rc=$?"
cat $O/r2o_band_probe.jsonl
cat $O/r2o_dyn_probe.jsonl
cat $O/r2o_prof_wf.log $O/r2o_prof_pool.log $O/r2o_prof_stream.log
ls -la $O | tail -30
# summaries on the box (the reports may be too big to travel back)
mkdir -p /tmp/elf && (cd /tmp/elf && cuobjdump -xelf all $ROOT/ray-tracer_b200/librtb.so > /dev/null && nvdisasm -gi -c rtb_render_fast.sm_100a.cubin > /tmp/fast.sass)
S="python tools/ncu_summary.py"
$S $O/r2o_wf_traverse.ncu-rep /tmp/fast.sass wf_traverse_kernelILb0ELb0 "wf_traverse_kernel<false,false>, C3 scene 1920x1080 (traversal launch 46 of a 4-spp run, one row group), round 2 final state" > $O/r2_ncu_wf_traverse_C3.txt 2>&1
$S $O/r2o_wf_shade.ncu-rep /tmp/fast.sass wf_shade_kernelILb0ELb0 "wf_shade_kernel<false,false>, C3 scene 1920x1080 (shade launch 46 of a 4-spp run, one row group), round 2 final state" > $O/r2_ncu_wf_shade_C3.txt 2>&1
$S $O/r2o_pool.ncu-rep /tmp/fast.sass pool_render_kernelILb0ELb0 "pool_render_kernel<false,false> (option kernel = 3), C3 scene 1920x1080, one 2-spp frame = one launch" > $O/r2_ncu_pool_C3.txt 2>&1
$S $O/r2o_stream.ncu-rep /tmp/fast.sass stream_render_kernelILb0ELb0 "stream_render_kernel<false,false> (option kernel = 4), C3 scene 1920x1080, one 2-spp frame = one launch" > $O/r2_ncu_stream_C3.txt 2>&1
head -30 $O/r2_ncu_wf_traverse_C3.txt
du -sh $O/*.ncu-rep
# keep the merge under its size limit: drop the biggest reports if the directory is too large
if [ $(du -sm $O | cut -f1) -gt 55 ]; then rm -f $O/r2o_pool.ncu-rep $O/r2o_stream.ncu-rep; fi
if [ $(du -sm $O | cut -f1) -gt 55 ]; then rm -f $O/r2o_wf_shade.ncu-rep; fi
