#!/bin/bash
# Round 2, last session: the five short GPU calls (4.3 GPU-minutes were left; each is `gpurun --timeout T -- '<command>'`).
# J: first hardware run of the Fortran-vector test + the 42 _maf cases after the residual fix       -> profiles/r02_f90c_gpu.log
#   timeout 180 python -m pytest tests/test_gpu_f90c_vectors.py tests/test_gpu_maf.py -q -m gpu
# K: smoke() and the whole GPU suite on the final product code                                    -> profiles/r02_smoke.log, r02_gpu_tests.log
#   timeout 40 python -c "import __graft_entry__ as g; g.smoke()"; timeout 170 python -m pytest tests -m gpu -x -q
# L: bench line with the reference's translated psor2sma_core + OpenMP as the CPU arm               -> profiles/r02_bench_line_refarm.json
#   timeout 108 python bench.py --steps 20 --warmup 3 --no-also --no-ttc
# M: CPU time-to-convergence leg with the reference's kernels on the box's 16 host threads          -> profiles/r02_cpu_ttc_reference.json
#   OMP_NUM_THREADS=$(nproc) OMP_PROC_BIND=close OMP_DYNAMIC=FALSE CUDA_VISIBLE_DEVICES="" timeout 72 python bench.py --cpu-worker ttc
# N: smoke() on the final snapshot
#   timeout 40 python -c "import __graft_entry__ as g; g.smoke()"
