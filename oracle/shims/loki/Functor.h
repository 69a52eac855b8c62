// TEST INFRASTRUCTURE ONLY: Loki::Functor / TL::MakeTypelist as far as the reference names them
// (PS_BlobTree/include/BlobTreeBuilder.h:52, a typedef that nothing on the polygonization path calls). See Singleton.h.
#ifndef PSGPU_ORACLE_LOKI_FUNCTOR_H
#define PSGPU_ORACLE_LOKI_FUNCTOR_H
#include <functional>
namespace Loki {
class NullType {};
namespace TL {
template <class... Ts> struct TypelistImpl {};
template <class... Ts> struct MakeTypelist { typedef TypelistImpl<Ts...> Result; };
}
template <class R, class TList> class Functor;
template <class R, class... Args>
class Functor<R, TL::MakeTypelist<Args...> > {
public:
    Functor() {}
    template <class F> Functor(F f) : m_f(f) {}
    R operator()(Args... a) { return m_f(a...); }
private:
    std::function<R(Args...)> m_f;
};
}
#endif
