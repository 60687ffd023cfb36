#pragma once
#define GTSAM_EXPORT
#define GTSAM_EXTERN_EXPORT extern
