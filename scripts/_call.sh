timeout 900 python -m pytest tests/test_dropin_integration.py -m gpu -q -k monotone 2>&1 | tail -30
