/* Host shim for firmware/src/util/cyclecounter.h (TEST INFRASTRUCTURE). */
#pragma once
void cycProfStart(const char *profile_name);
void cycProfEvent(const char *event);
void cycProfStop(void);
