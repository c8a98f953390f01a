/* TEST INFRASTRUCTURE ONLY: declarations-only stand-in for MATLAB's mex.h / matrix.h (MATLAB is not installed in this
 * image), just enough for `g++ -fsyntax-only` of a MEX gateway written against mexplus ("compile-check only
 * against a stub mex.h").  Signatures follow the documented MATLAB C Matrix API; nothing here is implemented. */
#ifndef ANISO_MEX_STUB_H
#define ANISO_MEX_STUB_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef struct mxArray_tag mxArray;
typedef size_t mwSize;
typedef size_t mwIndex;
typedef ptrdiff_t mwSignedIndex;
typedef bool mxLogical;
typedef uint16_t mxChar;
typedef enum { mxREAL, mxCOMPLEX } mxComplexity;
typedef enum {
    mxUNKNOWN_CLASS = 0, mxCELL_CLASS, mxSTRUCT_CLASS, mxLOGICAL_CLASS, mxCHAR_CLASS, mxVOID_CLASS, mxDOUBLE_CLASS,
    mxSINGLE_CLASS, mxINT8_CLASS, mxUINT8_CLASS, mxINT16_CLASS, mxUINT16_CLASS, mxINT32_CLASS, mxUINT32_CLASS,
    mxINT64_CLASS, mxUINT64_CLASS, mxFUNCTION_CLASS
} mxClassID;
void mexFunction(int nlhs, mxArray* plhs[], int nrhs, const mxArray* prhs[]);
void mexErrMsgIdAndTxt(const char* id, const char* fmt, ...);
void mexWarnMsgIdAndTxt(const char* id, const char* fmt, ...);
void mexErrMsgTxt(const char* msg);
void mexLock(void);
void mexUnlock(void);
int mexPrintf(const char* fmt, ...);
mxArray* mxCreateNumericMatrix(mwSize m, mwSize n, mxClassID cls, mxComplexity c);
mxArray* mxCreateNumericArray(mwSize ndim, const mwSize* dims, mxClassID cls, mxComplexity c);
mxArray* mxCreateDoubleMatrix(mwSize m, mwSize n, mxComplexity c);
mxArray* mxCreateDoubleScalar(double v);
mxArray* mxCreateLogicalMatrix(mwSize m, mwSize n);
mxArray* mxCreateLogicalScalar(bool v);
mxArray* mxCreateString(const char* s);
mxArray* mxCreateCharArray(mwSize ndim, const mwSize* dims);
mxArray* mxCreateCellMatrix(mwSize m, mwSize n);
mxArray* mxCreateStructMatrix(mwSize m, mwSize n, int nfields, const char** names);
mxArray* mxDuplicateArray(const mxArray* a);
void mxDestroyArray(mxArray* a);
void* mxGetData(const mxArray* a);
double* mxGetPr(const mxArray* a);
double mxGetScalar(const mxArray* a);
mxChar* mxGetChars(const mxArray* a);
mxLogical* mxGetLogicals(const mxArray* a);
int mxGetString(const mxArray* a, char* buf, mwSize len);
size_t mxGetNumberOfElements(const mxArray* a);
mwSize mxGetNumberOfDimensions(const mxArray* a);
const mwSize* mxGetDimensions(const mxArray* a);
size_t mxGetM(const mxArray* a);
size_t mxGetN(const mxArray* a);
mwSize mxGetNzmax(const mxArray* a);
mxClassID mxGetClassID(const mxArray* a);
const char* mxGetClassName(const mxArray* a);
mxArray* mxGetCell(const mxArray* a, mwIndex i);
void mxSetCell(mxArray* a, mwIndex i, mxArray* v);
mxArray* mxGetField(const mxArray* a, mwIndex i, const char* name);
void mxSetField(mxArray* a, mwIndex i, const char* name, mxArray* v);
int mxAddField(mxArray* a, const char* name);
int mxGetNumberOfFields(const mxArray* a);
const char* mxGetFieldNameByNumber(const mxArray* a, int n);
mwIndex mxCalcSingleSubscript(const mxArray* a, mwSize nsubs, const mwIndex* subs);
bool mxIsCell(const mxArray* a);
bool mxIsStruct(const mxArray* a);
bool mxIsChar(const mxArray* a);
bool mxIsEmpty(const mxArray* a);
bool mxIsSparse(const mxArray* a);
bool mxIsSingle(const mxArray* a);
bool mxIsDouble(const mxArray* a);
bool mxIsNumeric(const mxArray* a);
bool mxIsLogical(const mxArray* a);
bool mxIsLogicalScalar(const mxArray* a);
bool mxIsLogicalScalarTrue(const mxArray* a);
bool mxIsComplex(const mxArray* a);
bool mxIsClass(const mxArray* a, const char* name);
bool mxIsFromGlobalWS(const mxArray* a);
bool mxIsInt8(const mxArray* a);
bool mxIsInt16(const mxArray* a);
bool mxIsInt32(const mxArray* a);
bool mxIsInt64(const mxArray* a);
bool mxIsUint8(const mxArray* a);
bool mxIsUint16(const mxArray* a);
bool mxIsUint32(const mxArray* a);
bool mxIsUint64(const mxArray* a);
bool mxIsFunctionHandle(const mxArray* a);
bool mxIsNaN(double v);
bool mxIsInf(double v);
bool mxIsFinite(double v);
double mxGetNaN(void);
double mxGetInf(void);
double mxGetEps(void);
#ifdef __cplusplus
}
#endif
#endif
