"""A/B timing helper: run a scripts/*.py timing script against an alternative build of the library.
usage: python scripts/ab/run_ab.py <lib.so> <script.py> [args...]  (development tool, not product code)"""
import runpy, sys
sys.path.insert(0, '.')
from detectron2_centernet_b200 import _lib
_lib.LIB_PATH = sys.argv[1]
sys.argv = sys.argv[2:]
runpy.run_path(sys.argv[0], run_name='__main__')
