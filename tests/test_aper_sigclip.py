"""aper(skyalgorithm='sigmaclipping') (aper.py:275-282) against goldens from the unmodified reference."""
import warnings

import numpy as np
import pytest

from tests.conftest import assert_close, load_golden

CASES = {"exact_4050": dict(apr=np.array([3.0, 5.0, 8.0]), skyrad=[40, 50], badpix=[-12000, 60000], exact=True),
         "approx_1220_nobadpix": dict(apr=np.array([2.5, 6.0]), skyrad=[12, 20], badpix=[0, 0], exact=False)}


@pytest.mark.parametrize("name", sorted(CASES))
def test_sigclip_oracle_matches_reference_golden(name):
    from oracle import sigclip_oracle
    g, a = load_golden("golden_aper_sigclip.npz"), load_golden("golden_aper.npz")
    kw = CASES[name]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        sky, sig, nsky = sigclip_oracle.sky(a["image"].astype(np.float64), a["x"], a["y"], kw["skyrad"], kw["badpix"])
    assert_close(sky, g[name + "__sky"], rtol=1e-12, what="sky")
    assert_close(sig, g[name + "__skyerr"], rtol=1e-12, what="skyerr")


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_aper_sigclip_gpu_matches_reference_golden(name):
    from pythonphot_b200 import aper
    g, a = load_golden("golden_aper_sigclip.npz"), load_golden("golden_aper.npz")
    r = aper.aper(a["image"].astype(np.float64), a["x"], a["y"], phpadu=float(a["phpadu"]), zeropoint=25.0, verbose=False,
                  skyalgorithm="sigmaclipping", **CASES[name])
    n = a["x"].size
    # apertures whose flag hinges on the rounding of one atan (a NaN pixel with overlap weight 0 +- 1e-9) are
    # identified by the oracle from the geometry alone and left out, as in tests/test_gpu_parity.py
    from oracle import oracle
    fragile = oracle.aper(a["image"].astype(np.float64), a["x"], a["y"], phpadu=float(a["phpadu"]), zeropoint=25.0,
                          return_extra=True, **CASES[name])[9]
    ok = ~fragile
    assert np.array_equal(np.asarray(r[6]).reshape(n, -1)[ok], g[name + "__badflag"][ok])
    for nm, got in zip(("mag", "magerr", "flux", "fluxerr", "sky", "skyerr"), r[:6]):
        want = g["%s__%s" % (name, nm)]
        got = np.asarray(got).reshape(want.shape)
        if want.ndim == 2:
            got, want = got[ok], want[ok]
        assert_close(got, want, rtol=1e-9, atol=1e-7 if nm in ("flux", "fluxerr") else 0.0, what=nm)


@pytest.mark.gpu
def test_aper_sigclip_gpu_matches_oracle_sky_on_a_clean_field():
    """No NaN pixels: every star gets a clipped sky; 3000 stars, fp32 frame, crowded enough that clipping matters."""
    from oracle import sigclip_oracle
    from pythonphot_b200 import aper, synth
    f = synth.make_field(700, 800, 3000, seed=51, dtype=np.float32, margin=2.0, edge_frac=0.03, nan_frac=0.0, sat_frac=2e-4, with_noise=False)
    r = aper.aper(f["image"], f["x"], f["y"], phpadu=f["phpadu"], apr=np.array([4.0]), skyrad=[20, 30], badpix=[-12000, 60000], exact=False,
                  verbose=False, skyalgorithm="sigmaclipping", return_info=True)
    sky, sig, nsky = sigclip_oracle.sky(f["image"].astype(np.float64), f["x"], f["y"], [20, 30], [-12000, 60000])
    assert np.array_equal(r[8]["nsky"], nsky)
    assert_close(r[4], sky, rtol=1e-9, what="sky")
    assert_close(r[5], sig, rtol=1e-9, what="skyerr")
    assert (r[8]["mmm_iter"] > 1).mean() > 0.5  # the clip does remove stars from most annuli
