/* Hand-written stand-in for the autoconf-generated config.h of the reference
 * (SWIFT 0.9.0, PlanetSim fork).  TEST INFRASTRUCTURE ONLY: used by
 * oracle/build_ref.sh to compile the reference's own sources, where they lie
 * under /root/reference, into oracle/_ref/.  Never part of the product.
 *
 * Scheme selection: -DREF_PLANETARY picks PLANETARY_SPH + EOS_PLANETARY,
 * -DREF_SPHENIX picks SPHENIX_SPH + EOS_IDEAL_GAS (the reference's configure default,
 * configure.ac:1958-1959; it defines EXTRA_HYDRO_LOOP itself, src/part.h:84-90),
 * default is MINIMAL_SPH + EOS_IDEAL_GAS. */
#ifndef REF_CONFIG_H
#define REF_CONFIG_H

#ifndef _GNU_SOURCE
#define _GNU_SOURCE 1
#endif

#ifdef REF_PLANETARY
#define PLANETARY_SPH 1
#define EOS_PLANETARY 1
#elif defined(REF_SPHENIX)
#define SPHENIX_SPH 1
#define EOS_IDEAL_GAS 1
#else
#define MINIMAL_SPH 1
#define EOS_IDEAL_GAS 1
#endif

#define NONE_MHD 1
#define CUBIC_SPLINE_KERNEL 1
#define HYDRO_DIMENSION_3D 3
#define HYDRO_GAMMA_5_3 (5. / 3.)
#define DEFAULT_GRAVITY 1
#define SELF_GRAVITY_MULTIPOLE_ORDER 4
#define STARS_NONE 1
#define BLACK_HOLES_NONE 1
#define SINK_NONE 1
#define COOLING_NONE 1
#define CHEMISTRY_NONE 1
#define TRACERS_NONE 1
#define ENTROPY_FLOOR_NONE 1
#define PRESSURE_FLOOR_NONE 1
#define STAR_FORMATION_NONE 1
#define FEEDBACK_NONE 1
#define RT_NONE 1
#define EXTRA_IO_NONE 1
#define EXTERNAL_POTENTIAL_NONE 1
#define NEUTRINO_NONE 1

#define HAVE_POSIX_MEMALIGN 1
#define HAVE_PTHREAD_BARRIERS 1
#define HAVE_SETAFFINITY 1
#define HAVE_CBRTF 1
#define HAVE_EXP10 1
#define HAVE_EXP10F 1
#define HAVE_SINCOS 1
#define HAVE_SINCOSF 1
#define HAVE_IMMINTRIN_H 1
#define HAVE_CLOCK_GETTIME 1
#define HAVE_UNISTD_H 1
#define HAVE_SYS_TIME_H 1
#define HAVE_TIME_H 1
#define STDC_HEADERS 1

#define SWIFT_RANDOM_SEED_XOR 0
#define SWIFT_CONFIG_FLAGS "oracle-handwritten-config"
#define GIT_CMD "git"
#define PACKAGE_URL "www.swiftsim.com"
#define PACKAGE_NAME "SWIFT"
#define PACKAGE "swift"
#define VERSION "0.9.0"

#endif
