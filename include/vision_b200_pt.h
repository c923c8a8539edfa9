/* vision_b200_pt.h — path-tracing part of the C ABI (included by vision_b200.h). */
#ifndef VISION_B200_PT_H
#define VISION_B200_PT_H
/* filled in below as the megakernel lands */
#endif
