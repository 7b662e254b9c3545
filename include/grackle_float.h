/* grackle-b200: FP64 build only (north_star: computation is FP64 as in the reference).
 * Replaces the generated header of the reference build (src/clib/Makefile:205-208). */
#ifndef GRACKLE_B200_FLOAT_H
#define GRACKLE_B200_FLOAT_H
#define GRACKLE_FLOAT_8
#endif
