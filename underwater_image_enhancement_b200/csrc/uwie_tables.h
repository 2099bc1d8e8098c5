// Host-side table builders and the NumPy percentile planner (see uwie_tables.cpp).
#pragma once
#include "uwie_color.h"

namespace uwie {

void build_color_tables(ColorTables* t);

// np.percentile(..., method='linear') on n values: result = lerp(sorted[k], sorted[k+1], gamma)
// (k+1 clamped to n-1).  gamma is the value NumPy computes in the data's float type.
struct PercentilePlan {
    long long k;
    double gamma;
};
PercentilePlan percentile_plan_f32(long long n, double pct);
PercentilePlan percentile_plan_f64(long long n, double pct);

}  // namespace uwie
