"""Low-storage explicit Runge-Kutta (Carpenter-Kennedy) coefficients.

Values exactly as written in /root/reference/include/LSERK4.hpp:16-27 (the
reference has rk4a and rk4b only; there is no rk4c table).
Stage idiom (/root/reference/src/advec1d/main.cpp:92-102):
    res = rk4a[s]*res + dt*RHS ; u += rk4b[s]*res
"""
numStages = 5

rk4a = (
    0.0,
    -567301805773.0 / 1357537059087.0,
    -2404267990393.0 / 2016746695238.0,
    -3550918686646.0 / 2091501179385.0,
    -1275806237668.0 / 842570457699.0,
)

rk4b = (
    1432997174477.0 / 9575080441755.0,
    5161836677717.0 / 13612068292357.0,
    1720146321549.0 / 2090206949498.0,
    3134564353537.0 / 4481467310338.0,
    2277821191437.0 / 14882151754819.0,
)
