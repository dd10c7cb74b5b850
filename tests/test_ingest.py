"""CPU: collector text parsers (the build-defined ingestion formats)."""
import numpy as np

from kube_gpu_stats_b200 import ingest

PMON = """# gpu        pid  type    sm   mem   enc   dec    fb   command
# Idx          #   C/G     %     %     %     %    MB   name
    0      12345     C    37    12     -     -  4096   python
    0      12399     C     -     -     -     -    10   nvidia-cuda-mps
    1          -     -     -     -     -     -     -   -
    3       2211     C   100    55     0     0 70000   trainer
"""


def test_parse_pmon():
    gpu, pid, sm, mem, fb = ingest.parse_pmon(PMON)
    assert gpu.tolist() == [0, 0, 3] and pid.tolist() == [12345, 12399, 2211]
    assert sm.tolist() == [37, 0, 100] and mem.tolist() == [12, 0, 55] and fb.tolist() == [4096, 10, 65535]
    util = ingest.pack_util(sm, mem, fb)
    assert util.dtype == np.uint32 and util[0] == 37 | (12 << 8) | (4096 << 16)


def test_pod_of_cgroup_layouts():
    cgroupfs = ("12:devices:/kubepods/burstable/pod0a1b2c3d-1111-2222-3333-444455556666/"
                "9f8e7d\\n1:name=systemd:/kubepods/burstable/pod0a1b2c3d-1111-2222-3333-444455556666/9f8e7d\\n").replace('\\\\n', '\\n')
    systemd = "0::/kubepods.slice/kubepods-besteffort.slice/kubepods-besteffort-pod0A1B2C3D_1111_2222_3333_444455556666.slice/cri-containerd-abc.scope\\n"
    assert ingest.pod_of_cgroup(cgroupfs) == '0a1b2c3d-1111-2222-3333-444455556666'
    assert ingest.pod_of_cgroup(systemd) == '0a1b2c3d-1111-2222-3333-444455556666'
    assert ingest.pod_of_cgroup("0::/user.slice/user-1000.slice/session-3.scope\\n") is None


def test_series_major_order():
    ts = np.array([5, 5, 5, 6, 6, 6, 5], dtype=np.uint32)
    gid = np.array([1, 1, 0, 1, 1, 0, 0], dtype=np.uint32)
    pid = np.array([9, 7, 3, 9, 7, 3, 2], dtype=np.uint32)
    o = ingest.series_major_order(ts, gid, pid)
    assert list(zip(gid[o], pid[o], ts[o])) == [(0, 2, 5), (0, 3, 5), (0, 3, 6), (1, 7, 5), (1, 7, 6), (1, 9, 5), (1, 9, 6)]
