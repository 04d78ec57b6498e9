python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -5
for r in 2 4 8; do MG_JACOBI_ROWS=$r MG_RBGS_ROWS=$r MG_RESIDUAL_ROWS=$r python profiles/sweep_stages.py --level 12 --tag R$r; done
python profiles/sweep_stages.py --level 11 --tag auto
python profiles/profile_population.py
