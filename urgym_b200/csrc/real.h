// Scalar type of the build: float (product, liburgym_b200.so) or double (parity build, -DUG_F64, liburgym_b200_f64.so).
#pragma once
#ifdef UG_F64
typedef double real;
#else
typedef float real;
#endif
