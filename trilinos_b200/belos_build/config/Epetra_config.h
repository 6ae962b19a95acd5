/* Hand-written Epetra configuration for the oracle build (oracle/_ref): serial comm, no
 * Fortran kernels (C++ loops of Epetra_CrsMatrix.cpp:3352-3363,3516-3534), BLAS/LAPACK with
 * trailing-underscore symbols. */
#ifndef B200_EPETRA_CONFIG_H
#define B200_EPETRA_CONFIG_H
#ifndef F77_BLAS_MANGLE
#define F77_BLAS_MANGLE(name, NAME) name##_
#endif
#ifndef F77_FUNC
#define F77_FUNC(name, NAME) name##_
#endif
#ifndef F77_FUNC_
#define F77_FUNC_(name, NAME) name##_
#endif
#define HAVE_BLAS
#define HAVE_LAPACK
#define EPETRA_ADDRESS64BIT
#define HAVE_EPETRA_TEUCHOS
#ifndef EPETRA_DEPRECATED
#define EPETRA_DEPRECATED __attribute__((__deprecated__))
#endif
#ifndef EPETRA_DEPRECATED_MSG
#define EPETRA_DEPRECATED_MSG(MSG) __attribute__((__deprecated__))
#endif
#endif
