python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -5
for r in 2 4 8; do MG_JACOBI_ROWS=$r MG_RBGS_ROWS=$r MG_RESIDUAL_ROWS=$r python profiles/sweep_stages.py --level 12 --tag R$r; done
for l in 8 9 10 11; do python profiles/sweep_stages.py --level $l --tag auto; done
python bench.py --steps 1 --warmup 3
