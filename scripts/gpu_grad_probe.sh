mkdir -p gpurun_out
(python scripts/grad_probe2.py 600; python scripts/grad_probe2.py 0) > gpurun_out/grad_probe2.log 2>&1
grep -v Warning gpurun_out/grad_probe2.log
