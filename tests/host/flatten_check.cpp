// Host-only check of ProjectionFlatten (orbx.hpp): the values the reference's expressions give, computed independently here.
#include <cmath>
#include <cstdio>
#include <vector>
#include "orbx.hpp"
int main()
{
    std::vector<float> sf(8);
    sf[0] = 1.f;
    for (int i = 1; i < 8; ++i) sf[i] = (float)(sf[i - 1] * (double)1.2f);
    int bad = 0;
    auto expect = [&](bool ok, const char* what) { if (!ok) { std::printf("FAIL %s\n", what); ++bad; } };
    // 0.998f as a double is 0.99800002...: the reference's `viewCos > 0.998` is TRUE for it
    orbm_proj_query q = orbx::ProjectionFlatten::LocalMap(10.f, 20.f, 5.f, 2, 0.998f, 1.0f, sf);
    expect(q.radius == 2.5f * sf[2] && q.min_level == 1 && q.max_level == 2 && q.max_right_err == q.radius && q.u_right == 5.f, "localmap 0.998f");
    q = orbx::ProjectionFlatten::LocalMap(10.f, 20.f, 5.f, 0, 0.9f, 3.0f, sf);
    expect(q.radius == (4.0f * 3.0f) * sf[0] && q.min_level == -1 && q.max_level == 0, "localmap th 3");
    q = orbx::ProjectionFlatten::LastFrame(100.f, 50.f, 0.5f, 3, 45.f, 15.f, 40.f, false, false, sf);
    expect(q.radius == 15.f * sf[3] && q.min_level == 2 && q.max_level == 4 && q.u_right == 100.f - 40.f * 0.5f && q.angle == 45.f, "lastframe");
    q = orbx::ProjectionFlatten::LastFrame(100.f, 50.f, 0.5f, 3, 45.f, 15.f, 40.f, true, false, sf);
    expect(q.min_level == 3 && q.max_level == -1, "lastframe forward");
    q = orbx::ProjectionFlatten::LastFrame(100.f, 50.f, 0.5f, 3, 45.f, 15.f, 40.f, false, true, sf);
    expect(q.min_level == 0 && q.max_level == 3, "lastframe backward");
    q = orbx::ProjectionFlatten::KeyFrame(1.f, 2.f, 0, 7.f, 10.f, sf);
    expect(q.min_level == -1 && q.max_level == 1 && q.level_lo == -1 && q.radius == 10.f, "keyframe");
    q = orbx::ProjectionFlatten::Sim3OrFuse(1.f, 2.f, 1.f, 0, 10.f, 40.f, sf);
    expect(q.level_lo == 0 && q.level_hi == 0 && q.min_level == -1 && q.max_level == -1 && q.u_right == 1.f - 40.f, "sim3 level 0");
    q = orbx::ProjectionFlatten::Sim3OrFuse(1.f, 2.f, 1.f, 5, 3.f, 0.f, sf);
    expect(q.level_lo == 4 && q.level_hi == 5 && q.radius == 3.f * sf[5], "sim3 level 5");
    std::printf(bad ? "flatten check FAILED\n" : "flatten check ok\n");
    return bad;
}
