"""NumPy restatement of the non-affine scaler transforms the reference applies to the constituents (TEST INFRASTRUCTURE, see
oracle/__init__.py):   fit_scaler / apply_scaling   OE-VAE/utils.py:389-417
  QuantileTransformer(output_distribution='normal', n_quantiles=10000, random_state=0).transform   (utils.py:394, vae.sh:9-10)
  PowerTransformer().transform   (utils.py:396; yeo-johnson + standardize)
The arithmetic lives in scikit-learn 1.9.0 / SciPy 1.18.1 / NumPy 2.3 (third-party, unpinned by the reference):
  sklearn/preprocessing/_data.py  QuantileTransformer._transform_col, PowerTransformer.transform, StandardScaler.transform
  numpy/_core/src/multiarray/compiled_base.c  arr_interp;   scipy.special.ndtri (Cephes ndtri.c);   scipy.stats._morestats._yeojohnson_transform
restated element by element (this is what the CUDA kernels follow) and PINNED against the libraries themselves in tests/test_oracle.py.
"""
import numpy as np

BOUNDS_THRESHOLD = 1e-7

# ---- Cephes ndtri.c (inverse of the normal CDF), the algorithm behind scipy.special.ndtri / stats.norm.ppf
S2PI = 2.50662827463100050242E0
EXPM2 = 0.13533528323661269189
P0 = [-5.99633501014107895267E1, 9.80010754185999661536E1, -5.66762857469070293439E1, 1.39312609387279679503E1, -1.23916583867381258016E0]
Q0 = [1.95448858338141759834E0, 4.67627912898881538453E0, 8.63602421390890590575E1, -2.25462687854119370527E2, 2.00260212380060660359E2,
      -8.20372256168333339912E1, 1.59056225126211695515E1, -1.18331621121330003142E0]
P1 = [4.05544892305962419923E0, 3.15251094599893866154E1, 5.71628192246421288162E1, 4.40805073893200834700E1, 1.46849561928858024014E1,
      2.18663306850790267539E0, -1.40256079171354495875E-1, -3.50424626827848203418E-2, -8.57456785154685413611E-4]
Q1 = [1.57799883256466749731E1, 4.53907635128879210584E1, 4.13172038254672030440E1, 1.50425385692907503408E1, 2.50464946208309415979E0,
      -1.42182922854787788574E-1, -3.80806407691578277194E-2, -9.33259480895457427372E-4]
P2 = [3.23774891776946035970E0, 6.91522889068984211695E0, 3.93881025292474443415E0, 1.33303460815807542389E0, 2.01485389549179081538E-1,
      1.23716634817820021358E-2, 3.01581553508235416007E-4, 2.65806974686737550832E-6, 6.23974539184983293730E-9]
Q2 = [6.02427039364742014255E0, 3.67983563856160859403E0, 1.37702099489081330271E0, 2.16236993594496635890E-1, 1.34204006088543189037E-2,
      3.28014464682127739104E-4, 2.89247864745380683936E-6, 6.79019408009981274425E-9]


def _polevl(x, c):
    a = np.full_like(x, c[0])
    for k in c[1:]: a = a * x + k
    return a


def _p1evl(x, c):
    a = x + c[0]
    for k in c[1:]: a = a * x + k
    return a


def ndtri(y0):
    """Cephes ndtri in float64, vectorised with the same operation order."""
    y0 = np.asarray(y0, np.float64)
    out = np.full(y0.shape, np.nan)
    with np.errstate(all='ignore'):
        out[y0 == 0.0] = -np.inf; out[y0 == 1.0] = np.inf
        ok = (y0 > 0.0) & (y0 < 1.0)
        y = y0.copy(); code = np.ones(y0.shape, bool)
        flip = ok & (y > 1.0 - EXPM2)
        y[flip] = 1.0 - y[flip]; code[flip] = False
        mid = ok & (y > EXPM2)
        ym = y[mid] - 0.5; y2 = ym * ym
        out[mid] = (ym + ym * (y2 * _polevl(y2, P0) / _p1evl(y2, Q0))) * S2PI
        tail = ok & ~mid
        yt = y[tail]
        x = np.sqrt(-2.0 * np.log(yt)); x0 = x - np.log(x) / x; z = 1.0 / x
        x1 = np.where(x < 8.0, z * _polevl(z, P1) / _p1evl(z, Q1), z * _polevl(z, P2) / _p1evl(z, Q2))
        r = x0 - x1
        out[tail] = np.where(code[tail], -r, r)
    return out


def interp(x, xp, fp):
    """numpy.interp (arr_interp in compiled_base.c) for float64 inputs: j = last index with xp[j] <= x; exact hits return fp[j];
    slope * (x - xp[j]) + fp[j] with slope = (fp[j+1] - fp[j]) / (xp[j+1] - xp[j]); fp[0] / fp[-1] outside; NaN stays NaN."""
    x = np.asarray(x, np.float64); xp = np.asarray(xp, np.float64); fp = np.asarray(fp, np.float64)
    n = len(xp)
    out = np.empty(x.shape)
    nan = np.isnan(x)
    j = np.clip(np.searchsorted(xp, np.where(nan, xp[0], x), side='right') - 1, 0, n - 1)
    jn = np.minimum(j + 1, n - 1)
    with np.errstate(all='ignore'):
        slope = (fp[jn] - fp[j]) / (xp[jn] - xp[j])
        val = slope * (x - xp[j]) + fp[j]
        bad = np.isnan(val)                                              # "if we get nan in one direction, try the other"
        alt = slope * (x - xp[jn]) + fp[jn]
        val = np.where(bad, np.where(np.isnan(alt) & (fp[j] == fp[jn]), fp[j], alt), val)
    out[:] = val
    out[(j == n - 1) | (xp[j] == x)] = fp[j][(j == n - 1) | (xp[j] == x)]
    out[x > xp[-1]] = fp[-1]; out[x < xp[0]] = fp[0]
    out[nan] = np.nan
    return out


def quantile_normal_transform(X, quantiles, references):
    """QuantileTransformer(output_distribution='normal')._transform_col on every column of a float32 matrix X [N, D].
    quantiles [n_q, D] float64, references [n_q] float64.  Rounding sequence for a float32 input (sklearn operates in place on the
    float32 array): bound tests on float32(x -+ 1e-7); 0.5 (interp - interp) in float64 -> float32; bounds -> 1 / 0; ndtri of that
    float32 value (SciPy's float32 loop: float32 result); clip in float64; -> float32."""
    X = np.asarray(X, np.float32)
    out = np.empty_like(X)
    clip_min = float(ndtri(np.array([BOUNDS_THRESHOLD - np.spacing(1)]))[0])
    clip_max = float(ndtri(np.array([1 - (BOUNDS_THRESHOLD - np.spacing(1))]))[0])
    for f in range(X.shape[1]):
        x = X[:, f]
        q = quantiles[:, f]
        with np.errstate(invalid='ignore'):
            lower = (x - np.float32(BOUNDS_THRESHOLD)).astype(np.float64) < q[0]
            upper = (x + np.float32(BOUNDS_THRESHOLD)).astype(np.float64) > q[-1]
        xd = x.astype(np.float64)
        t = (0.5 * (interp(xd, q, references) - interp(-xd, -q[::-1], -references[::-1]))).astype(np.float32)
        t[upper] = 1; t[lower] = 0
        v = ndtri(t.astype(np.float64)).astype(np.float32)
        out[:, f] = np.clip(v.astype(np.float64), clip_min, clip_max).astype(np.float32)
    return out


def power_transform(X, lambdas, mean=None, scale=None):
    """PowerTransformer(method='yeo-johnson', standardize=True).transform on a float32 matrix.  A transformer fitted on float32 data
    holds float32 lambdas_, so SciPy's _yeojohnson_transform runs entirely in float32: expm1(lmbda * log1p(x)) / lmbda for x >= 0,
    -expm1((2 - lmbda) * log1p(-x)) / (2 - lmbda) for x < 0 (log1p alone when |lmbda| resp. |lmbda - 2| < float32 eps); then
    StandardScaler.transform in place with mean_ / scale_ cast to float32: (y - mean) / scale in float32."""
    X = np.asarray(X, np.float32)
    out = np.empty_like(X)
    eps = np.finfo(np.float32).eps
    for f in range(X.shape[1]):
        x, lm = X[:, f], np.float32(lambdas[f])
        pos = x >= 0
        y = np.zeros(x.shape, np.float32)
        with np.errstate(all='ignore'):
            y[pos] = np.log1p(x[pos]) if abs(lm) < eps else np.expm1(lm * np.log1p(x[pos])) / lm
            y[~pos] = -np.expm1((2 - lm) * np.log1p(-x[~pos])) / (2 - lm) if abs(lm - 2) > eps else -np.log1p(-x[~pos])
        out[:, f] = y
    if mean is not None:
        out = out - np.asarray(mean, np.float32)
    if scale is not None:
        out = out / np.asarray(scale, np.float32)
    return out
