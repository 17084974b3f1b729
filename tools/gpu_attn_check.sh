timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -8 | cut -c1-300
python tools/microbench.py 2>/dev/null | grep enc_attn
WB200_ATTN_TWOPASS=1 python tools/microbench.py 2>/dev/null | grep enc_attn
