import sys
sys.path.insert(0, '.')
import __graft_entry__ as g
pkg = g.load_package()
from oracle import oracle
from tests import test_gpu_queue as T
T._drive(pkg, oracle, NF=6, n_imu=120, frame_every=8, delay=0.0302, check_each=True)
