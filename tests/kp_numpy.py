"""Second, independent restatement (pure Python / numpy) of the reference's key-point rules, used only to
cross-check the C++ oracle's integer outputs.  Reference: src/KeyPointGenerator/KeyPointGenerator.cpp
  set_interval     :319-339      adaptive (jerk)   :341-382      velocity_change :642-728
  jerk profile     :730-770      velocity profile  :797-808      percentages     :814-838
  interpolation    :840-954
X is (T+1, 2n): [pos(n) ; vel(n)] per time index (n == dof_quat in every in-scope config).
"""
import numpy as np


def set_interval(T, n, min_N):
    full = list(range(n))
    rows = [full[:] if t % min_N == 0 else [] for t in range(T - 1)]
    rows.append(full[:])
    return rows


def jerk_profile(X, n, T, dt):
    jerk = np.zeros((T, n))
    for t in range(T - 2):
        a1 = (X[t + 1] - X[t]) / dt
        a2 = (X[t + 2] - X[t + 1]) / dt
        for j in range(n):
            jerk[t, j] = abs((a2[j + n] - a1[j + n]) / dt)
    jerk[T - 2] = 0.0
    jerk[T - 1] = 0.0
    return jerk


def adaptive_jerk(X, n, T, dt, min_N, max_N, thr):
    prof = jerk_profile(X, n, T, dt)
    rows = [list(range(n))]
    last = [0] * n
    for t in range(1, T - 1):
        row = []
        for j in range(n):
            if t - last[j] >= min_N:
                if prof[t, j] > thr[j]:
                    row.append(j)
                    last[j] = t
            if t - last[j] >= max_N:
                row.append(j)
                last[j] = t
        rows.append(row)
    rows.append(list(range(n)))
    return rows


def velocity_change(X, n, T, min_N, max_N, thr):
    vel = X[:, n:2 * n]
    rows = [list(range(n))]
    cnt = [0] * n
    acc = [0.0] * n
    last_dir = [0.0] * n
    for t in range(1, T):
        row = []
        for i in range(n):
            cnt[i] += 1
            cur_dir = vel[t, i] - vel[t - 1, i]
            acc[i] += abs(vel[t, i])
            if cnt[i] >= min_N and abs(acc[i]) > thr[i]:
                row.append(i)
                acc[i] = 0.0
                cnt[i] = 0
                continue
            if cnt[i] >= min_N:
                if cur_dir * last_dir[i] < 0:
                    row.append(i)
                    acc[i] = 0.0
                    cnt[i] = 0
                    continue
            else:
                last_dir[i] = cur_dir
            if cnt[i] >= max_N:
                row.append(i)
                acc[i] = 0.0
                cnt[i] = 0
                continue
        rows.append(row)
    rows[T - 1].extend(range(n))
    return rows


def percentages(rows, n):
    T = len(rows)
    count = [0] * n
    for r in rows:
        for v in r:
            count[v] += 1
    return np.array([c / T * 100 for c in count])


def interpolate(rows, n, m, A, B):
    """A (T, nx, nx), B (T, nx, m) stored [t][col][row] (column-major blocks); in place.
    col[k] = col[s] + (k - s) * ((col[e] - col[s]) / (e - s))  (KeyPointGenerator.cpp:905-944)."""
    T = len(rows)
    for i in range(n):
        s = 0
        for t in range(1, T):
            if i not in rows[t]:
                continue
            e = t
            span = float(e - s)
            for cols, M in (((i, i + n), A), ((i,) if i < m else (), B)):
                for c in cols:
                    add = (M[e, c] - M[s, c]) / span
                    for k in range(s + 1, e):
                        M[k, c] = M[s, c] + float(k - s) * add
            s = e
