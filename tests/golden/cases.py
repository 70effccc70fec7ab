"""Case table shared by tests/golden/gen_golden.py and the parity tests
(reference keyword names, aper.py:26-29)."""
import numpy as np

APER_CASES = [
    # name, kwargs (reference keyword names, aper.py:26-29)
    ("exact_4050", dict(apr=np.array([3.0, 5.0, 8.0]), skyrad=[40, 50], badpix=[-12000, 60000], exact=True)),
    ("approx_4050", dict(apr=np.array([3.0, 5.0, 8.0]), skyrad=[40, 50], badpix=[-12000, 60000], exact=False)),
    ("exact_1220_nobadpix", dict(apr=np.array([2.5, 6.0]), skyrad=[12, 20], badpix=[0, 0], exact=True)),
    ("exact_single_apr", dict(apr=5.0, skyrad=[40, 50], badpix=[-12000, 60000], exact=True)),
    ("exact_setsky", dict(apr=np.array([3.0, 5.0]), skyrad=[40, 50], badpix=[-12000, 60000], exact=True,
                          setskyval=[300.0, 11.0, 500.0])),
    ("approx_setsky_scalar", dict(apr=np.array([4.0]), skyrad=[40, 50], badpix=[-12000, 60000], exact=False,
                                  setskyval=295.0)),
    ("exact_readnoise", dict(apr=np.array([4.0]), skyrad=[15, 25], badpix=[-12000, 60000], exact=True,
                             readnoise=5.0)),
    ("exact_minsky_big", dict(apr=np.array([3.0]), skyrad=[5, 7], badpix=[-12000, 60000], exact=True, minsky=70)),
]
