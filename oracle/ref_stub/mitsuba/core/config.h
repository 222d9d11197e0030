// Stub for the generated mitsuba/core/config.h (the real one is produced by
// mitsuba3/resources/configure.py:17-105).  Lets the reference's header-only math compile in
// scalar float mode for oracle/_ref.  Contains no reference code.
#pragma once
#define MI_VERSION_MAJOR 3
#define MI_VERSION_MINOR 6
#define MI_VERSION_PATCH 0
#define MI_VARIANTS "scalar_rgb\n"
#define MI_DEFAULT_VARIANT "scalar_rgb"
#define MI_INSTANTIATE_CLASS(Name)
#define MI_INSTANTIATE_STRUCT(Name)
#define MI_EXTERN_CLASS(Name)
#define MI_EXTERN_STRUCT(Name)
#define MI_VARIANT_TEMPLATE()
