"""Host/device coherence for lattice configurations.

The reference keeps a configuration as a float64 numpy array inside a small holder object
(``examples/ising2D.py:56-62``) that models and observers read and write freely.  Here the live
copy is in HBM; ``LatticeConfig`` keeps the same attributes (``lenX``, ``lenY``, ``config``) and
moves data only when a side that does not own the newest copy asks for it:

* ``config`` (read)  -> download if the device copy is newer; because the caller gets a writable
  array, the device copy is then treated as out of date (it is re-uploaded at the next run).
* ``peek()``         -> read-only view, keeps the device copy valid (used by our own observers).
* engine start       -> upload if the host copy is the newest.
"""
import numpy as np

from . import _cabi
from .seeding import get_seed


def _layout_code(layout, lenY):
    if layout in ("auto", None):
        return _cabi.LAYOUT_BITS if lenY % 128 == 0 else _cabi.LAYOUT_INT8
    if layout in ("int8", _cabi.LAYOUT_INT8):
        return _cabi.LAYOUT_INT8
    if layout in ("bits", "bitpacked", _cabi.LAYOUT_BITS):
        return _cabi.LAYOUT_BITS
    raise ValueError("unknown layout %r (use 'auto', 'int8' or 'bits')" % (layout,))


class LatticeConfig:
    """Base of the drop-in config classes (``ising2D_config``, ``latgas_config``)."""

    _pmc_values = (-1, 1)  # host values of the two site states (low, high)

    def __init__(self, lenX, lenY, layout="auto", dtype=np.float64, device=0):
        self.lenX = int(lenX)
        self.lenY = int(lenY)
        self.layout = layout
        self.device = device
        self._host = np.empty([self.lenX, self.lenY], dtype=dtype)
        self._host_valid = True
        self._dev_valid = False
        self._lat = None        # _cabi.Lattices holding this configuration ...
        self._rep = 0           # ... as local replica number _rep
        self._lat_shared = False  # True when the lattice batch belongs to a replica-exchange engine
        self._lat_key = None

    # -- host side ---------------------------------------------------------------------
    @property
    def config(self):
        self._pull()
        self._dev_valid = False  # the caller may write into the array
        return self._host

    @config.setter
    def config(self, value):
        self._host = np.asarray(value)
        self._host_valid = True
        self._dev_valid = False

    def peek(self):
        """Read-only view of the current configuration; does not invalidate the device copy."""
        self._pull()
        v = self._host.view()
        v.flags.writeable = False
        return v

    def _pull(self):
        if not self._host_valid:
            img = self._lat.download(self._rep)
            self._host[...] = self._from_int8(img)
            self._host_valid = True

    def _to_int8(self, a):
        if getattr(a, "dtype", None) == np.int8:
            return a
        lo, hi = self._pmc_values
        return np.where(np.asarray(a) > (lo + hi) / 2.0, hi, lo).astype(np.int8)

    def _from_int8(self, img):
        return img.astype(self._host.dtype)

    # -- device side -------------------------------------------------------------------
    def _attach(self, model_code):
        """Return the one-replica device lattice for this config, creating/uploading as needed."""
        layout = _layout_code(self.layout, self.lenY)
        key = (model_code, layout)
        if self._lat is not None and self._lat_shared:
            # bound to a batch of a replica-exchange engine: a host edit goes back into that batch, whatever layout
            # the batch was created with (it is taken from the first configuration)
            if self._lat_key[0] != model_code:
                raise TypeError("this configuration is bound to a batch of another model")
            if not self._dev_valid:
                self._lat.upload(self._rep, self._to_int8(self._host))
                self._dev_valid = True
            return self._lat
        if self._lat is None or self._lat_key != key:
            if self._lat is not None:
                self._pull()
                if not self._lat_shared:
                    self._lat.close()
                self._dev_valid = False
                self._lat_shared = False
                self._rep = 0
            self._lat = _cabi.Lattices(model_code, self.lenX, self.lenY, nreplicas=1, ntemps=1, seed=get_seed(),
                                       layout=layout, device=self.device)
            self._lat_key = key
        if not self._dev_valid:
            self._lat.upload(self._rep, self._to_int8(self._host))
            self._dev_valid = True
        return self._lat

    def _device_changed(self):
        self._host_valid = False
        self._dev_valid = True

    def _device_observables(self):
        """(bond, second) exact integers from the device if it holds the newest copy, else None."""
        if self._lat is not None and self._dev_valid:
            b, s = self._lat.tracked()
            return int(b[self._rep]), int(s[self._rep])
        return None

    def _bind(self, lat, rep, model_code):
        """Make this config local replica ``rep`` of a shared batch (replica exchange); uploads it."""
        if self._lat is not None and not self._lat_shared and self._lat is not lat:
            self._pull()
            self._lat.close()
        self._pull()
        self._lat, self._rep, self._lat_shared = lat, rep, True
        self._lat_key = (model_code, lat.layout)
        lat.upload(rep, self._to_int8(self._host))
        self._dev_valid = True

    # -- pickling (TemperatureRX_MPI / ParallelMC pickle configs, py_mc/mc_mpi.py:67,179) -------
    def __getstate__(self):
        self._pull()
        st = dict(self.__dict__)
        st["_lat"] = None
        st["_rep"] = 0
        st["_lat_shared"] = False
        st["_lat_key"] = None
        st["_dev_valid"] = False
        st["_host_valid"] = True
        return st

    def __deepcopy__(self, memo):
        import copy
        new = self.__class__.__new__(self.__class__)
        for k, v in self.__getstate__().items():
            setattr(new, k, copy.deepcopy(v, memo))
        return new
