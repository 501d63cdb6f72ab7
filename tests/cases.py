"""Shared case tables and input patterns for the parity tests (same generators as
tests/golden/make_golden.py)."""
import importlib.util
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_spec = importlib.util.spec_from_file_location("make_golden", os.path.join(_HERE, "golden", "make_golden.py"))
make_golden = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(make_golden)

SMALL_CASES = make_golden.SMALL_CASES
DIGEST_CASES = make_golden.DIGEST_CASES
PLAN_CASES = make_golden.PLAN_CASES
make_input = make_golden.make_input
