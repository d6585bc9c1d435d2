/* jni_min.h -- NOT a JNI implementation.  The few declarations of <jni.h> that systemml_b200_jni.cpp uses, so that the
 * generated forwarders can be type-checked against include/sysml_b200.h (g++ -fsyntax-only -DSYSML_JNI_SYNTAX_CHECK) in
 * an image that has no JDK.  A real build uses the JDK's <jni.h>; the shapes below follow the JNI specification. */
#ifndef SYSML_JNI_MIN_H
#define SYSML_JNI_MIN_H
#include <stdint.h>
typedef int32_t jint;
typedef int64_t jlong;
typedef double jdouble;
typedef int8_t jbyte;
typedef uint8_t jboolean;
typedef jint jsize;
class _jobject {};
class _jclass : public _jobject {};
class _jstring : public _jobject {};
class _jarray : public _jobject {};
class _jlongArray : public _jarray {};
class _jintArray : public _jarray {};
class _jdoubleArray : public _jarray {};
class _jbyteArray : public _jarray {};
typedef _jobject* jobject;
typedef _jclass* jclass;
typedef _jstring* jstring;
typedef _jarray* jarray;
typedef _jlongArray* jlongArray;
typedef _jintArray* jintArray;
typedef _jdoubleArray* jdoubleArray;
typedef _jbyteArray* jbyteArray;
#define JNI_ABORT 2
#define JNIEXPORT __attribute__((visibility("default")))
#define JNICALL
struct JNIEnv {
  void* GetPrimitiveArrayCritical(jarray array, jboolean* isCopy);
  void ReleasePrimitiveArrayCritical(jarray array, void* carray, jint mode);
  void SetLongArrayRegion(jlongArray array, jsize start, jsize len, const jlong* buf);
  const char* GetStringUTFChars(jstring str, jboolean* isCopy);
  void ReleaseStringUTFChars(jstring str, const char* chars);
  jstring NewStringUTF(const char* utf);
};
#endif
