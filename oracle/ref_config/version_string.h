#ifndef SWIFT_VERSION_STRING_H
#define SWIFT_VERSION_STRING_H
#define PACKAGE_VERSION "0.9.0"
#define GIT_REVISION "f745f98c"
#define GIT_BRANCH "subtask_dev"
#define GIT_DATE "unknown"
#define SWIFT_CFLAGS "-O2 -fno-fast-math -ffp-contract=off"
#endif
