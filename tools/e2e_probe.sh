set -e
mkdir -p gpurun_out/e2e
for kind in cpu gpu; do
  d=/tmp/e2e_$kind; rm -rf $d; mkdir -p $d; cp tests/golden/e2e/OPTIONS tests/golden/e2e/corot7.txt $d/
  (cd $d && $GRAFT_REPO_ROOT/host/_build/kima_${kind}_CoRoT7 -t 4 -s 43 > out.txt 2>&1)
  cp $d/sample.txt gpurun_out/e2e/CoRoT7_$kind.txt
done
cmp gpurun_out/e2e/CoRoT7_cpu.txt gpurun_out/e2e/CoRoT7_gpu.txt && echo "CPU == GPU on this box"
cmp gpurun_out/e2e/CoRoT7_cpu.txt tests/golden/e2e/CoRoT7_sample_seed43.txt && echo "CPU == golden" || echo "CPU != golden"
lscpu | grep "Model name"
