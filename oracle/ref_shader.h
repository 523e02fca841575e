/* ref_shader.h -- plain-float interface of ref_shader.cpp (TEST INFRASTRUCTURE ONLY). */
#ifndef REF_SHADER_H
#define REF_SHADER_H
#include "vr_oracle.h"
/* Bind the uniform block, volume and decoded LUT (not re-entrant: one render at a time). */
void ref_shader_bind(const vro_frame *f, const vro_volume *vol, const float *lut_rgba, uint32_t nlut);
/* Run volume.frag main() for one fragment; returns the number of volume fetches. */
uint32_t ref_shader_run(const float tex[3], const float frag[3], float out[4]);
#endif
