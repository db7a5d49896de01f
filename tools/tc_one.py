import sys
sys.path.insert(0, ".")
from tools.tc_check import run
M, K, N, p = [int(x) for x in sys.argv[1:5]]
run(M, K, N, p)
