"""Run the reference's UNMODIFIED main.py under Python >= 3.12.

`handyrl/evaluation.py:305` calls `random.randrange(1e8)`; Python 3.12 removed the acceptance of integral floats
there (TypeError), so `main.py --eval` of the reference dies on this image's interpreter (3.12.3) before it
touches any environment - with the reference's own envs too.  This launcher restores the <= 3.11 behaviour of
`random.randrange` for integral floats (an interpreter-compat shim, nothing of the reference is edited or
wrapped) and then executes main.py as `__main__` with the caller's argv and cwd.

    python tools/ref_main_py312.py baseline/_ref/main.py --eval models/1.pth 20 2
"""
import random
import runpy
import sys


def _compat_randrange():
    orig = random.Random.randrange

    def _int(x):
        if isinstance(x, float) and x == int(x):
            return int(x)
        return x

    def randrange(self, start, stop=None, step=1):
        return orig(self, _int(start), _int(stop), _int(step))

    random.Random.randrange = randrange
    random.randrange = random._inst.randrange


if __name__ == "__main__":
    main_py = sys.argv[1]
    sys.argv = [main_py] + sys.argv[2:]
    import os
    sys.path.insert(0, os.path.dirname(os.path.abspath(main_py)))
    _compat_randrange()
    runpy.run_path(main_py, run_name="__main__")
