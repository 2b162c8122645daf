// TEST INFRASTRUCTURE ONLY.  ps/cluster_region.h:31 includes the generated header of
// ps/detected_objects.proto but uses nothing from it; the export path is out of scope.
#ifndef ORACLE_SHIMS_PS_DETECTED_OBJECTS_PB_H_
#define ORACLE_SHIMS_PS_DETECTED_OBJECTS_PB_H_
#endif
