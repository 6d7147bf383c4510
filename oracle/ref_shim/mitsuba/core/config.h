/* Hand-written stand-in for the header the reference's build generates (resources/configure.py) -- TEST
 * INFRASTRUCTURE.  It declares one variant, scalar_rgb, and leaves the explicit-instantiation macros empty, which
 * is all the header-only templates compiled by oracle/ref_harness.cpp need.  Nothing is copied from the reference. */
#pragma once
#include <mitsuba/core/fwd.h>
#define MTS_VARIANTS "scalar_rgb\n"
#define MTS_DEFAULT_VARIANT "scalar_rgb"
#define MTS_EXTERN_STRUCT_CORE(Name)
#define MTS_EXTERN_CLASS_CORE(Name)
#define MTS_EXTERN_STRUCT_RENDER(Name)
#define MTS_EXTERN_CLASS_RENDER(Name)
#define MTS_INSTANTIATE_STRUCT(Name)
#define MTS_INSTANTIATE_CLASS(Name)
#define MTS_INVOKE_VARIANT(variant, func, ...) func<float, mitsuba::Color<float, 3>>(__VA_ARGS__)
