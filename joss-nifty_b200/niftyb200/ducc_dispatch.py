"""Drop-in for ``nifty8.ducc_dispatch`` (reference ``src/ducc_dispatch.py``).

Same module-level names: ``fftn``, ``ifftn``, ``hartley``, ``vdot``, ``set_nthreads``,
``nthreads``.  :func:`install` rebinds them inside an imported ``nifty8`` (the three
modules that ``from``-import them: SURVEY.md section 8b), so that every
``HartleyOperator`` / ``FFTOperator`` / ``Field.vdot`` of the reference runs on the GPU.
"""

from .transforms import fftn, hartley, ifftn, vdot  # noqa: F401

_nthreads = 1


def nthreads():
    """``ducc_dispatch.nthreads`` (``src/ducc_dispatch.py:29-30``)."""
    return _nthreads


def set_nthreads(nthr):
    """``ducc_dispatch.set_nthreads`` (``:33-35``).  Accepted for compatibility; the GPU
    backend ignores it (one CUDA stream per call)."""
    global _nthreads
    _nthreads = int(nthr)


def install(nifty8=None, sync_convention=True):
    """Patch an imported ``nifty8`` so that its ducc_dispatch seam calls this backend.

    ``harmonic_operators.py:27`` and ``field.py:24`` bind the functions with
    ``from .. import``, so the names have to be replaced in those modules too."""
    import importlib
    from . import config as _cfg
    if nifty8 is None:
        nifty8 = importlib.import_module("nifty8")
    dd = importlib.import_module(nifty8.__name__ + ".ducc_dispatch")
    ho = importlib.import_module(nifty8.__name__ + ".operators.harmonic_operators")
    fld = importlib.import_module(nifty8.__name__ + ".field")
    ncfg = importlib.import_module(nifty8.__name__ + ".config")

    def _hartley(a, axes=None):
        if sync_convention:
            _cfg.update("hartley_convention", ncfg._config.get("hartley_convention"))
        return hartley(a, axes=axes)

    for mod, names in ((dd, ("fftn", "ifftn", "hartley", "vdot")), (ho, ("fftn", "ifftn", "hartley")),
                       (fld, ("vdot",))):
        for n in names:
            setattr(mod, n, _hartley if n == "hartley" else globals()[n])
    return nifty8
