export PYTHONPATH=$PWD
for sp in 1; do timeout 300 python scripts/prof_attn.py --trace --splits $sp 2>&1 | grep -v "^  [a-z_]* *[0-9]" | tail -15; done
