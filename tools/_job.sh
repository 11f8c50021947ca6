set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r02d_pytest.txt; cat gpurun_out/r02d_pytest.txt
python tools/ktrace.py --configs cfg1,cfg2 --tp 1 > gpurun_out/r02d_ktrace_tp1.jsonl 2> gpurun_out/r02d_ktrace.err
python tools/ktrace.py --configs cfg5,cfg4 --tp 8 >> gpurun_out/r02d_ktrace_tp8.jsonl 2>> gpurun_out/r02d_ktrace.err
tail -3 gpurun_out/r02d_ktrace.err
for v in tree no_cvt no_mma no_ldmatrix no_exp only_copy; do
  if [ $v = tree ]; then unset B200PA_PKG_ROOT; else export B200PA_PKG_ROOT=_ab/$v; fi
  B200PA_AB_LABEL=$v python tools/powerprobe.py --config cfg4 --seconds 4 >> gpurun_out/r02d_powerprobe.jsonl 2>> gpurun_out/r02d_powerprobe.err
  sleep 3
done
unset B200PA_PKG_ROOT
cut -c1-420 gpurun_out/r02d_powerprobe.jsonl
tail -3 gpurun_out/r02d_powerprobe.err
