set -x
timeout 1500 python scripts/kslice_probe.py > gpurun_out/r14_kslice.md 2> gpurun_out/r14_kslice.err
cat gpurun_out/r14_kslice.md; tail -3 gpurun_out/r14_kslice.err
