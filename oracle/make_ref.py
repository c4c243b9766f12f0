"""Recipe for ``oracle/_ref/`` (test / measurement infrastructure; the directory is git-ignored and never
committed, it travels to the GPU box with the working tree like the built ``.so`` files).

The reference is a directory of Python scripts without packaging, so "building" it is placing the two files its
FOM path needs where ``bench.py``'s CPU legs can import them on a machine that has no ``/root/reference``:

    BurgersFD_CleanFine/hypernet2D.py   (the module the drop-in replaces; imported UNMODIFIED)
    BurgersFD_CleanFine/rbf_utils.py    (imported by it at module level)

plus nothing else: ``matplotlib`` is not installed in this image and the module touches ``plt.rcParams`` at
import time; ``load()`` puts the same two-attribute stand-in into ``sys.modules`` that
``oracle/gen_golden.py`` uses.  ``__graft_entry__.build()`` runs ``make()`` when ``/root/reference`` is present.

    python oracle/make_ref.py
"""
import contextlib
import io
import os
import shutil
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/BurgersFD_CleanFine"
REF_DST = os.path.join(HERE, "_ref")
FILES = ("hypernet2D.py", "rbf_utils.py")


def make():
    """Copy the reference's files into oracle/_ref/ (no-op without /root/reference). Returns the directory or None."""
    if not all(os.path.exists(os.path.join(REF_SRC, f)) for f in FILES):
        return REF_DST if available() else None
    os.makedirs(REF_DST, exist_ok=True)
    for f in FILES:
        src, dst = os.path.join(REF_SRC, f), os.path.join(REF_DST, f)
        if not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(src):
            shutil.copyfile(src, dst)
    return REF_DST


def available():
    return all(os.path.exists(os.path.join(REF_DST, f)) for f in FILES)


def load():
    """The reference module, unmodified, imported from oracle/_ref/ under the name it has in the reference's own
    scripts (``hypernet2D``); None when the directory has not been made."""
    if not available():
        return None
    if "matplotlib" not in sys.modules:
        class _RC(dict):
            def update(self, *a, **k):
                pass
        mpl = types.ModuleType("matplotlib")
        plt = types.ModuleType("matplotlib.pyplot")
        plt.rcParams = _RC()
        plt.rc = lambda *a, **k: None
        mpl.pyplot = plt
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = plt
    if REF_DST not in sys.path:
        sys.path.insert(0, REF_DST)
    with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
        import hypernet2D                       # noqa: E402  (the reference's file)
    if os.path.dirname(os.path.abspath(hypernet2D.__file__)) != REF_DST:
        return None                             # some other hypernet2D was imported first
    return hypernet2D


if __name__ == "__main__":
    print(make())
