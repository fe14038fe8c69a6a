/* TEST INFRASTRUCTURE ONLY: stubs for HDF5-side symbols of the reference that the hydro path never reaches. Filled from the link report. */
#include <stdio.h>
#include <stdlib.h>
#define REF_STUB(name) \
  void name(void) { fprintf(stderr, "oracle/_ref: HDF5-side stub %s reached\n", #name); abort(); }
REF_STUB(lightcone_array_trigger_map_update)
REF_STUB(do_line_of_sight)
REF_STUB(neutrino_response_struct_dump)
REF_STUB(neutrino_response_struct_restore)
REF_STUB(io_is_double_precision)
REF_STUB(los_struct_restore)
REF_STUB(los_struct_dump)
REF_STUB(lightcone_array_struct_restore)
REF_STUB(lightcone_array_struct_dump)
