"""tests-side entry to the oracle binding (oracle/oracle_py.py) — TEST INFRASTRUCTURE."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
from oracle_py import ORACLE_HIT_DTYPE, ORACLE_PATH, OracleScene, camera_eval, lib  # noqa: E402,F401
import oracle_py  # noqa: E402,F401
