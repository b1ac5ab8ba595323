import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built_lib():
    from coral_b200 import build
    return build.build()


_DATASETS = {}


@pytest.fixture(scope="session")
def dataset(tmp_path_factory):
    """dataset(name, **kw) -> dict(data=SimData, paths={bam,cn_seg,seed}); cached per session."""
    from sim import configs
    from sim.generate import simulate, write_dataset

    def get(name, **kw):
        key = (name, repr(sorted(kw.items())))
        if key not in _DATASETS:
            cfg = configs.CONFIGS[name](**kw)
            data = simulate(cfg)
            out = tmp_path_factory.mktemp("ds_%s" % name)
            paths = write_dataset(data, str(out))
            _DATASETS[key] = {"data": data, "paths": paths, "cfg": cfg}
        return _DATASETS[key]
    return get
