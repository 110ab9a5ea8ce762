"""reference python/hetu/_base.py: native-library loading helpers (``_LIB``, ``check_call``,
``c_array``, the DNNL availability table).  Libraries here are loaded lazily by ``_lib.py``."""
import ctypes

from . import _lib as _L
from . import cpu_links as _cpu_links

DNNL_LIB = {name: True for name in dir(_cpu_links) if not name.startswith('_')}


def _load_lib():
    return _L.cuda_lib() or _L.cpu_lib()


class _LazyLib(object):
    def __getattr__(self, name):
        lib = _load_lib()
        if lib is None:
            raise RuntimeError('native libraries are not built; run `python -c "import __graft_entry__ as g; g.build()"`')
        return getattr(lib, name)


_LIB = _LazyLib()


def check_call(ret):
    assert ret == 0, 'native call failed with code %s' % ret


def c_array(ctype, values):
    return (ctype * len(values))(*values)
