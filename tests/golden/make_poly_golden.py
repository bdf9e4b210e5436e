"""Extracts the reference's polynomial coefficient model (data/networks/fitted_models_casadi.pkl) into
tests/golden/poly_model.npz, without CasADi.

The pickle holds, per coefficient, the sklearn LinearRegression / PolynomialFeatures fit
('fitted_models': coef[34], intercept, degree-3 features of (alpha, beta, aileron, elevator); script
main/surrogates/polynomial_model.py:228-271) next to CasADi Functions generated from exactly those
numbers (create_casadi_function, polynomial_model.py:76-104).  The CasADi objects are replaced by stubs
while unpickling; the numeric content is copied unchanged.

Run once in the build container:  python tests/golden/make_poly_golden.py
"""
import pickle
import warnings
from pathlib import Path

import numpy as np

SRC = Path("/root/reference/data/networks/fitted_models_casadi.pkl")
OUT = Path(__file__).resolve().parent / "poly_model.npz"


class _Stub:
    def __init__(self, *a, **k):
        pass

    def __setstate__(self, st):
        self.state = st


class _Unpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if module.startswith("casadi"):
            return type(name, (_Stub,), {"__module__": module})
        return super().find_class(module, name)


def main():
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        d = _Unpickler(open(SRC, "rb")).load()
    fm = d["fitted_models"]
    names = ["CX", "CY", "CZ", "Cl", "Cm", "Cn"]
    assert list(fm) == names and list(d["casadi_functions"]) == names
    powers = np.asarray(fm["CX"]["poly"].powers_, np.int64)
    for k in names:
        assert np.array_equal(np.asarray(fm[k]["poly"].powers_), powers)
        assert np.array_equal(np.ravel(fm[k]["model"].coef_), fm[k]["coef"])
        assert float(fm[k]["model"].intercept_) == float(fm[k]["intercept"])
    coef = np.stack([np.asarray(fm[k]["coef"], np.float64) for k in names])
    intercept = np.array([float(fm[k]["intercept"]) for k in names])
    np.savez(OUT, powers=powers, coef=coef, intercept=intercept, features=np.array(["alpha", "beta", "aileron", "elevator"]))
    print("wrote", OUT, powers.shape, coef.shape)
    # the linear model's 6x6 table (q, alpha, beta, aileron, elevator, intercept), LinearModel coefficient_models.py:80-84
    lin = np.loadtxt("/root/reference/data/networks/linearised.csv", delimiter=",", skiprows=1)
    assert lin.shape == (6, 6)
    np.savez(OUT.parent / "linear_model.npz", table=lin, columns=np.array(["q", "alpha", "beta", "aileron", "elevator", "intercept"]))
    print("wrote", OUT.parent / "linear_model.npz")


if __name__ == "__main__":
    main()
