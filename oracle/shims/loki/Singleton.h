// TEST INFRASTRUCTURE ONLY: the few names of Loki's <loki/Singleton.h> the reference uses
// (PS_FrameWork/include/PS_ErrorManager.h:8,74; PS_BlobTree/include/BlobTreeBuilder.h:20-44). Loki is a third-party dependency
// of the reference that is neither vendored nor version-pinned (Parsip.pro points at ../Dependencies/LOKI); it only owns object
// lifetimes and registries here, no arithmetic. Written from Loki's documented interface (Alexandrescu, Modern C++ Design ch. 6).
#ifndef PSGPU_ORACLE_LOKI_SINGLETON_H
#define PSGPU_ORACLE_LOKI_SINGLETON_H
namespace Loki {
template <class T> struct CreateUsingNew { static T* Create() { return new T; } static void Destroy(T* p) { delete p; } };
template <class T> struct DefaultLifetime {};
template <class T> struct PhoenixSingleton {};
template <class T> struct NoDestroy {};
namespace LongevityLifetime {
template <class T> struct DieAsSmallObjectChild {};
template <class T> struct DieAsSmallObjectParent {};
}
template <class T, template <class> class CreationPolicy = CreateUsingNew, template <class> class LifetimePolicy = DefaultLifetime>
class SingletonHolder {
public:
    typedef T ObjectType;
    static T& Instance() {
        static T* inst = CreationPolicy<T>::Create();  // never destroyed: the registries must outlive every static user
        return *inst;
    }
private:
    SingletonHolder();
};
}
#endif
