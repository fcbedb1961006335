// oracle/shim/hdf5_hl.h  --  TEST INFRASTRUCTURE: the reference includes it (include/h5utils.h:5) but uses nothing from it.
#pragma once
