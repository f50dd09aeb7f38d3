"""Minimal stand-in for the pytest-asyncio plugin (not installed in this image, no network): runs
``@pytest.mark.asyncio`` coroutine tests with ``asyncio.run`` so V-Gate's own test-suite
(tests/conftest.py:27 declares ``pytest_plugins = ['pytest_asyncio']``) can run unmodified as the
regression gate for "the plumbing works unchanged with the b200 backend registered"."""
import asyncio
import inspect

import pytest

fixture = pytest.fixture


def pytest_configure(config):
    config.addinivalue_line("markers", "asyncio: run the coroutine test in an event loop")


@pytest.hookimpl(tryfirst=True)
def pytest_pyfunc_call(pyfuncitem):
    fn = pyfuncitem.obj
    if inspect.iscoroutinefunction(fn):
        args = {a: pyfuncitem.funcargs[a] for a in pyfuncitem._fixtureinfo.argnames}
        asyncio.run(fn(**args))
        return True
    return None
