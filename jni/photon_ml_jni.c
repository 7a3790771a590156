/* JNI stub binding include/photon_ml_b200.h -- NOT compiled in this image (no JVM / jni.h).
   Kept in sync with INTEGRATION.md section 2. */
#include <jni.h>
#include "photon_ml_b200.h"

static void throw_for(JNIEnv *env, int rc) {
    const char *cls = rc == PML_EINVAL ? "java/lang/IllegalArgumentException" : "java/lang/RuntimeException";
    (*env)->ThrowNew(env, (*env)->FindClass(env, cls), pml_last_error());
}

JNIEXPORT jlong JNICALL Java_com_linkedin_photon_ml_function_glm_NativeGLM_ctxCreate(JNIEnv *env, jclass c, jint device) {
    pml_ctx *ctx = NULL; int rc = pml_ctx_create(device, &ctx);
    if (rc) { throw_for(env, rc); return 0; }
    return (jlong)(intptr_t)ctx;
}

JNIEXPORT jlong JNICALL Java_com_linkedin_photon_ml_function_glm_NativeGLM_datasetCreate(
        JNIEnv *env, jclass c, jlong ctx, jlong n, jint d, jlongArray rowPtr, jintArray col, jdoubleArray val,
        jdoubleArray labels, jdoubleArray offsets, jdoubleArray weights) {
    /* GetPrimitiveArrayCritical on each array (NULL-safe for offsets/weights), then: */
    pml_dataset *ds = NULL;
    int rc = pml_dataset_create_csr((pml_ctx *)(intptr_t)ctx, n, d, rp, ci, v, PML_F64, PML_STORE_AUTO, y, o, w, &ds);
    /* ReleasePrimitiveArrayCritical(..., JNI_ABORT) on each */
    if (rc) { throw_for(env, rc); return 0; }
    return (jlong)(intptr_t)ds;
}

JNIEXPORT jdouble JNICALL Java_com_linkedin_photon_ml_function_glm_NativeGLM_valueAndGradient(
        JNIEnv *env, jclass c, jlong ds, jint loss, jdoubleArray coef, jdoubleArray outGrad) {
    jdouble value = 0;
    double *w = (*env)->GetPrimitiveArrayCritical(env, coef, NULL);
    double *g = outGrad ? (*env)->GetPrimitiveArrayCritical(env, outGrad, NULL) : NULL;
    int rc = pml_value_and_gradient((pml_dataset *)(intptr_t)ds, loss, w, &value, g);
    if (g) (*env)->ReleasePrimitiveArrayCritical(env, outGrad, g, 0);
    (*env)->ReleasePrimitiveArrayCritical(env, coef, w, JNI_ABORT);
    if (rc) throw_for(env, rc);
    return value;
}
/* hessianVector(ds, loss, coef, tag, v, out), hessianDiagonal(ds, loss, coef, out), setOffsets, setWeights,
   setNormalization(ds, factors, shifts, interceptId), datasetDestroy, ctxDestroy: same pattern, one ABI call each. */
