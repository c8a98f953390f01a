// MEX gateway of lowrank/aniso rebuilt on libaniso_b200: the six ops of the reference's AnisoWrapper.cpp
// (AnisoWrapper.cpp:10-138: new, delete, getNodes, setCoeff, cache, mapping) with the same names and argument lists, so
// class/@Aniso/Aniso.m, aniso.m and demo.m run unchanged, plus two ops that replace the mapping loops of aniso.m:121-172.
// Build (on a machine with MATLAB):  mex -I<reference>/mexplus -I<repo>/include AnisoWrapper_b200.cpp -L<repo>/aniso_b200 -laniso_b200
// Here MATLAB is absent: `make -C oracle mexcheck` compiles this file with -fsyntax-only against oracle/mex_stub/mex.h and
// the reference's own mexplus headers.
#include <mexplus.h>

#include <map>
#include <memory>

#include "aniso_b200.h"

using namespace mexplus;

namespace {

void ck(int rc) {
    if (rc < 0) mexErrMsgIdAndTxt("aniso:b200", "%s", aniso_last_error());
}

// handle table: int64 keys like mexplus::Session (dispatch.h:183-275), but the opaque aniso_t is released with
// aniso_destroy instead of delete
std::map<int64_t, aniso_t*>& table() {
    static std::map<int64_t, aniso_t*> t;
    return t;
}
int64_t next_id = 0;

aniso_t* get(const mxArray* a) {
    const int64_t id = static_cast<int64_t>(mxGetScalar(a));
    std::map<int64_t, aniso_t*>::iterator it = table().find(id);
    if (it == table().end()) mexErrMsgIdAndTxt("mexplus:session:notFound", "Invalid id %d. Did you create?", (int)id);
    return it->second;
}

}  // namespace

// AnisoWrapper('new', ds, qr, ks, as, sr, fn, fm) -> int64 handle (AnisoWrapper.cpp:10-31; scalars arrive as doubles and are
// truncated with int(*p))
MEX_DEFINE(new)(int nlhs, mxArray* plhs[], int nrhs, const mxArray* prhs[]) {
    InputArguments in(nrhs, prhs, 7);
    OutputArguments out(nlhs, plhs, 1);
    aniso_t* h = nullptr;
    ck(aniso_create(int(in.get<double>(0)), int(in.get<double>(1)), int(in.get<double>(2)), in.get<double>(3),
                    int(in.get<double>(4)), int(in.get<double>(5)), int(in.get<double>(6)), &h));
    table()[++next_id] = h;
    mexLock();
    out.set(0, next_id);
}

MEX_DEFINE(delete)(int nlhs, mxArray* plhs[], int nrhs, const mxArray* prhs[]) {
    InputArguments in(nrhs, prhs, 1);
    OutputArguments out(nlhs, plhs, 0);
    const int64_t id = static_cast<int64_t>(mxGetScalar(prhs[0]));
    aniso_destroy(get(prhs[0]));
    table().erase(id);
    mexUnlock();
}

// N x 2 double, x column then y column (AnisoWrapper.cpp:33-44)
MEX_DEFINE(getNodes)(int nlhs, mxArray* plhs[], int nrhs, const mxArray* prhs[]) {
    InputArguments in(nrhs, prhs, 1);
    OutputArguments out(nlhs, plhs, 1);
    aniso_t* h = get(prhs[0]);
    const int n = aniso_num_nodes(h);
    plhs[0] = mxCreateNumericMatrix(n, 2, mxDOUBLE_CLASS, mxREAL);
    ck(aniso_get_nodes(h, mxGetPr(plhs[0]), mxGetPr(plhs[0]) + n, nullptr));
}

// setCoeff(h, sigma_s, sigma_t): interpolation; singPrecompute; makeKernels (AnisoWrapper.cpp:46-70)
MEX_DEFINE(setCoeff)(int nlhs, mxArray* plhs[], int nrhs, const mxArray* prhs[]) {
    InputArguments in(nrhs, prhs, 3);
    OutputArguments out(nlhs, plhs, 0);
    ck(aniso_set_coeff(get(prhs[0]), mxGetPr(prhs[1]), mxGetPr(prhs[2])));
}

// cache(h, id): the four cache steps (AnisoWrapper.cpp:72-90)
MEX_DEFINE(cache)(int nlhs, mxArray* plhs[], int nrhs, const mxArray* prhs[]) {
    InputArguments in(nrhs, prhs, 2);
    OutputArguments out(nlhs, plhs, 0);
    ck(aniso_cache(get(prhs[0]), int(*mxGetPr(prhs[1]))));
}

// mapping(h, charge, id) -> N x 1 (AnisoWrapper.cpp:92-136)
MEX_DEFINE(mapping)(int nlhs, mxArray* plhs[], int nrhs, const mxArray* prhs[]) {
    InputArguments in(nrhs, prhs, 3);
    OutputArguments out(nlhs, plhs, 1);
    aniso_t* h = get(prhs[0]);
    plhs[0] = mxCreateNumericMatrix(aniso_num_nodes(h), 1, mxDOUBLE_CLASS, mxREAL);
    ck(aniso_mapping(h, mxGetPr(prhs[1]), int(*mxGetPr(prhs[2])), mxGetPr(plhs[0])));
}

// blockForward(h, u, weighted) -> K*N x 1: aniso.m:121-136 (weighted = 0) / 138-156 (weighted = 1) in one call
MEX_DEFINE(blockForward)(int nlhs, mxArray* plhs[], int nrhs, const mxArray* prhs[]) {
    InputArguments in(nrhs, prhs, 3);
    OutputArguments out(nlhs, plhs, 1);
    aniso_t* h = get(prhs[0]);
    const int K = aniso_kernel_size(h) / 2 + 1;
    plhs[0] = mxCreateNumericMatrix((size_t)K * aniso_num_nodes(h), 1, mxDOUBLE_CLASS, mxREAL);
    ck(aniso_block_forward(h, mxGetPr(prhs[1]), int(*mxGetPr(prhs[2])), mxGetPr(plhs[0])));
}

// [y, flag, iter] = blockSolve(h, rhs, restart, tol, maxit): GMRES on I - mforward (aniso.m:158-172, no preconditioner)
MEX_DEFINE(blockSolve)(int nlhs, mxArray* plhs[], int nrhs, const mxArray* prhs[]) {
    InputArguments in(nrhs, prhs, 5);
    OutputArguments out(nlhs, plhs, 3);
    aniso_t* h = get(prhs[0]);
    const int K = aniso_kernel_size(h) / 2 + 1;
    plhs[0] = mxCreateNumericMatrix((size_t)K * aniso_num_nodes(h), 1, mxDOUBLE_CLASS, mxREAL);  // zero initial guess
    int iters = 0;
    const int rc = aniso_block_gmres(h, mxGetPr(prhs[1]), mxGetPr(plhs[0]), int(*mxGetPr(prhs[2])), int(*mxGetPr(prhs[4])),
                                     *mxGetPr(prhs[3]), &iters, nullptr, 0);
    ck(rc);
    if (nlhs > 1) plhs[1] = mxCreateDoubleScalar(rc);
    if (nlhs > 2) plhs[2] = mxCreateDoubleScalar(iters);
}

MEX_DISPATCH
