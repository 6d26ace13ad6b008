// nn_bvh_b200/csrc/spec_body.inl -- node phase and leaf phase of SpecKernel (included inside its
// outer loop; see spec.cuh for the state and the exactness argument).

        // ---- phase A: one step per lane per iteration ----
        while (true) {
            const bool busy = pend < 0;
            const bool isNode = cur >= 0;
            const bool isLeaf = cur < kCurDead;
            const bool canStep = !stall && (isNode ? (SPEC || !busy) : (isLeaf && !busy));
            const unsigned stepMask = __ballot_sync(0xffffffffu, canStep);
            const unsigned blockedMask = __ballot_sync(0xffffffffu, busy && !canStep);
            // idle lanes: blocked behind a postponed leaf, or (while rays remain) without a ray.
            // Live lanes are exactly stepMask | blockedMask.
            const unsigned idle = __popc(blockedMask) + (exhausted ? 0u : 32u - __popc(stepMask | blockedMask));
            if (stepMask == 0 || idle >= leafQuorum) break;
            if (canStep) {
                bool needPop = false;
                if (!(curT < tMax)) {
                    // the reference's `tMin < raytMax` (vecmath.h:1607) for the node / leaf the lane
                    // steps onto.  tLim == tMax unless a leaf is postponed.
                    if (SPEC && curT < tLim) stall = true;
                    else needPop = true;
                } else if (!isNode) {
                    // a leaf, and nothing postponed yet: postpone it and move on
                    pend = cur;
                    tLim = SpecLimit(tMax);
                    needPop = true;
                } else {
                    const unsigned char *np = wideBase + (size_t)(unsigned)cur * sizeof(WideNode);
                    const F8 a = Ldg256(np), b = Ldg256(np + 32);
                    // WideNode: lo0 xyz, hi0 xyz, lo1 xy | lo1 z, hi1 xyz, ref0, ref1, axis, pad
                    const bool nx = ry.neg & 1, ny = ry.neg & 2, nz = ry.neg & 4;
                    const float n0x = nx ? a.v[3] : a.v[0], f0x = nx ? a.v[0] : a.v[3];
                    const float n0y = ny ? a.v[4] : a.v[1], f0y = ny ? a.v[1] : a.v[4];
                    const float n0z = nz ? a.v[5] : a.v[2], f0z = nz ? a.v[2] : a.v[5];
                    const float n1x = nx ? b.v[1] : a.v[6], f1x = nx ? a.v[6] : b.v[1];
                    const float n1y = ny ? b.v[2] : a.v[7], f1y = ny ? a.v[7] : b.v[2];
                    const float n1z = nz ? b.v[3] : b.v[0], f1z = nz ? b.v[0] : b.v[3];
                    const int ref0 = __float_as_int(b.v[4]), ref1 = __float_as_int(b.v[5]);
                    const unsigned axis = __float_as_uint(b.v[6]);
                    const bool secondFirst = (ry.neg >> axis) & 1u;  // aggregates.cpp:561
                    const int refN = secondFirst ? ref1 : ref0, refF = secondFirst ? ref0 : ref1;
                    float tN, tF;
                    bool okN, okF;  // the tMax-independent part of the two box tests
                    if (EXACT) {
                        RayConst rc = ToRayConst(ry);
                        float t0, t1;
                        const bool ok0 = SlabStatic(rc, n0x, f0x, n0y, f0y, n0z, f0z, &t0);
                        const bool ok1 = SlabStatic(rc, n1x, f1x, n1y, f1y, n1z, f1z, &t1);
                        tN = secondFirst ? t1 : t0;
                        tF = secondFirst ? t0 : t1;
                        okN = secondFirst ? ok1 : ok0;
                        okF = secondFirst ? ok0 : ok1;
                    } else {
                        const float pad = SlabPad();
                        const float ax0 = Mul(Sub(n0x, ry.ox), ry.ix), bx0 = Mul(Mul(Sub(f0x, ry.ox), ry.ix), pad);
                        const float ay0 = Mul(Sub(n0y, ry.oy), ry.iy), by0 = Mul(Mul(Sub(f0y, ry.oy), ry.iy), pad);
                        const float az0 = Mul(Sub(n0z, ry.oz), ry.iz), bz0 = Mul(Mul(Sub(f0z, ry.oz), ry.iz), pad);
                        const float ax1 = Mul(Sub(n1x, ry.ox), ry.ix), bx1 = Mul(Mul(Sub(f1x, ry.ox), ry.ix), pad);
                        const float ay1 = Mul(Sub(n1y, ry.oy), ry.iy), by1 = Mul(Mul(Sub(f1y, ry.oy), ry.iy), pad);
                        const float az1 = Mul(Sub(n1z, ry.oz), ry.iz), bz1 = Mul(Mul(Sub(f1z, ry.oz), ry.iz), pad);
                        const float t0 = fmaxf(fmaxf(ax0, ay0), az0), x0 = fminf(fminf(bx0, by0), bz0);
                        const float t1 = fmaxf(fmaxf(ax1, ay1), az1), x1 = fminf(fminf(bx1, by1), bz1);
                        tN = secondFirst ? t1 : t0;
                        tF = secondFirst ? t0 : t1;
                        const float xN = secondFirst ? x1 : x0, xF = secondFirst ? x0 : x1;
                        okN = (tN <= xN) && (xN > 0.f);
                        okF = (tF <= xF) && (xF > 0.f);
                    }
                    const bool nearPass = okN && (tN < tMax);
                    const bool farPass = okF && (tF < tMax);
                    bool undecidable = false;
                    if (SPEC)  // empty window unless a leaf is postponed (then tLim > tMax)
                        undecidable = (okN && !nearPass && tN < tLim) ||
                                      (!nearPass && okF && !farPass && tF < tLim);
                    if (undecidable) {
                        stall = true;  // stay AT this node until the postponed leaf is resolved
                    } else if (nearPass) {
                        // The far child is stacked whenever its tMax-independent test passes: tMax
                        // can GROW by an ulp before the pop (shapes.cpp:239-247), see DESIGN.md.
                        if (okF) stack[sp++] = ((unsigned long long)__float_as_uint(tF) << 32) | (unsigned)refF;
                        cur = refN;
                        curT = tN;
                    } else if (farPass) {
                        cur = refF;
                        curT = tF;
                    } else {
                        needPop = true;
                    }
                }
                if (needPop) NNBVH_POP();
            }
        }

        // ---- phase B: every lane with a postponed leaf tests ONE triangle of it ----
        if (pend < 0) {
            const uint32_t prim = (uint32_t)pend & 0x7fffffffu;
            const float4 *tp = tris + 3 * (size_t)prim;
            const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
            const uint32_t flags = __float_as_uint(v0.w);
            bool done = (flags & kTriLast) != 0;
            if (!(flags & kTriDegenerate)) {
                RayConst rc = ToRayConst(ry);
                float t, b0, b1, b2;
                if (TriangleTest(rc, tMax, v0.x, v0.y, v0.z, v1.x, v1.y, v1.z, v2.x, v2.y, v2.z, &t,
                                 &b0, &b1, &b2)) {
                    if (ANY) {  // first accepted primitive ends the ray (:598-601)
                        hit.prim = 1u;
                        cur = kCurNoMore;
                        done = true;
                    } else if (SPEC && !(t <= tLim)) {
                        // grew beyond the bound the speculative decisions relied on: retrace
                        // this ray in the non-speculative exact pass
                        uint32_t pos = atomicAdd(p.nDeferred, 1u);
                        p.deferred[pos] = rayIdx;
                        rayIdx = kNoIdx;
                        cur = kCurNoMore;
                        done = true;
                    } else {
                        hit.prim = prim;
                        hit.t = t;
                        hit.b0 = b0;
                        hit.b1 = b1;
                        hit.b2 = b2;
                        tMax = t;
                        tLim = fminf(tLim, SpecLimit(t));
                    }
                }
            }
            if (done) {
                pend = 0;
                tLim = tMax;
                if (cur == kCurNoMore) cur = kCurDead;
            } else {
                pend = pend + 1;  // next primitive of the same leaf (bit 31 stays set)
            }
            stall = false;
        }
