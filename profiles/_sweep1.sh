for r in 8 16 24 32 45 64 91; do MG_JACOBI_ROWS=$r MG_RBGS_ROWS=$r MG_RESIDUAL_ROWS=$r MG_RESTRICT_ROWS=$r MG_PROLONG_ROWS=$r python profiles/sweep_stages.py --level 12 --tag rows$r; done
for r in 2 4 8 16 32; do MG_JACOBI_ROWS=$r MG_RBGS_ROWS=$r MG_RESIDUAL_ROWS=$r MG_RESTRICT_ROWS=$r MG_PROLONG_ROWS=$r python profiles/sweep_stages.py --level 10 --tag rows$r; done
