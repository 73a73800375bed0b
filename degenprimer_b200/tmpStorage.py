"""Temp-file hand-off of results, as BioUtils/Tools/tmpStorage.py:121-131 does.

`SearchEngine.find` returns the *path* of a temporary store holding
`(fwd_results, rev_results)` under the key 'result' (DegenPrimer/SearchEngine.py:385);
callers read it back with `from_shelf` (DegenPrimer/PCR_ProductsFinder.py:45-55).
The store is a pickle file instead of a dbm shelf (one object, read once);
files are removed at exit like the reference's registry does (:54-58).
"""
import atexit
import os
import pickle
import tempfile

_registry = set()


def _cleanup_all():
    for f in list(_registry):
        cleanup_file(f)


atexit.register(_cleanup_all)


def register_tmp_file(filename):
    _registry.add(filename)


def cleanup_file(filename):
    _registry.discard(filename)
    try:
        os.unlink(filename)
    except OSError:
        pass


def cleanup_files(filenames):
    for f in filenames:
        cleanup_file(f)


def to_shelf(result):
    if result is None:
        return None
    fd, name = tempfile.mkstemp(prefix='degenprimer_b200_', suffix='.result')
    with os.fdopen(fd, 'wb') as fh:
        pickle.dump({'result': result}, fh, protocol=pickle.HIGHEST_PROTOCOL)
    register_tmp_file(name)
    return name


def from_shelf(filename, cleanup=True):
    with open(filename, 'rb') as fh:
        return pickle.load(fh).get('result')
