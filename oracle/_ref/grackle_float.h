#ifndef __GRACKLE_FLOAT_H__
#define __GRACKLE_FLOAT_H__
#define GRACKLE_FLOAT_8
#endif
