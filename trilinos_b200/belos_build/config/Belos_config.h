/* Hand-written Belos configuration: timers on (the Teuchos::TimeMonitor breakdown is part of
 * the measurement plan), no Epetra / Tpetra / TSQR adapters in the product build. */
#ifndef B200_BELOS_CONFIG_H
#define B200_BELOS_CONFIG_H
#define HAVE_MUTABLE
#define HAVE_NAMESPACES
#define HAVE_NEW_FOR_SCOPING
#define BELOS_TEUCHOS_TIME_MONITOR
#ifdef B200_REF_BUILD
#define HAVE_BELOS_EPETRA
#endif
#ifndef BELOS_DEPRECATED
#define BELOS_DEPRECATED __attribute__((__deprecated__))
#endif
#ifndef BELOS_DEPRECATED_MSG
#define BELOS_DEPRECATED_MSG(MSG) __attribute__((__deprecated__))
#endif
#endif
