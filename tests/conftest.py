import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pk():
    """the product package (its directory name has a hyphen, hence importlib)"""
    return importlib.import_module("sphinxsys-learning_b200")


@pytest.fixture(scope="session")
def oracle():
    """CPU restatement of the reference: the checker, never the thing under test"""
    from oracle import binding
    binding.load()
    return binding


@pytest.fixture(scope="session")
def gpu(pk):
    """the CUDA backend module (host.py over libsphb200.so); fails loudly when absent"""
    pk.capi.load()
    return pk.host
