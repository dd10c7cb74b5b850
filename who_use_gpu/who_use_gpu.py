#!/usr/bin/env python3
"""Same path and zero-argument CLI as the reference script; the work is in kube_gpu_stats_b200."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kube_gpu_stats_b200.who_use_gpu import *  # noqa: F401,F403,E402
from kube_gpu_stats_b200.who_use_gpu import main  # noqa: E402

if __name__ == '__main__':
    main()
