// TEST INFRASTRUCTURE ONLY: Loki::Factory / Loki::CloneFactory with the members the reference calls: Register, Unregister,
// CreateObject (PS_BlobTree/include/C*.cpp register every node class by name, by BlobNodeType and by typeid; CBlobTree.cpp and
// the scene reader create / clone nodes through them). See Singleton.h.
#ifndef PSGPU_ORACLE_LOKI_FACTORY_H
#define PSGPU_ORACLE_LOKI_FACTORY_H
#include <map>
#include <stdexcept>
#include <typeinfo>
#include <typeindex>
#include "Singleton.h"
#include "Functor.h"
namespace Loki {
class TypeInfo {
public:
    TypeInfo(const std::type_info& ti) : m_ti(&ti) {}
    const std::type_info& Get() const { return *m_ti; }
    bool before(const TypeInfo& o) const { return m_ti->before(*o.m_ti) != 0; }
    const char* name() const { return m_ti->name(); }
private:
    const std::type_info* m_ti;
};
inline bool operator<(const TypeInfo& a, const TypeInfo& b) { return a.before(b); }
inline bool operator==(const TypeInfo& a, const TypeInfo& b) { return a.Get() == b.Get(); }

template <class AbstractProduct, typename IdentifierType, typename ProductCreator = AbstractProduct* (*)()>
class Factory {
public:
    bool Register(const IdentifierType& id, ProductCreator creator) { return m_map.insert(std::make_pair(id, creator)).second; }
    bool Unregister(const IdentifierType& id) { return m_map.erase(id) == 1; }
    AbstractProduct* CreateObject(const IdentifierType& id) {
        typename std::map<IdentifierType, ProductCreator>::iterator it = m_map.find(id);
        if (it == m_map.end()) throw std::runtime_error("Unknown Type");
        return (it->second)();
    }
private:
    std::map<IdentifierType, ProductCreator> m_map;
};

template <class AbstractProduct, class ProductCreator = AbstractProduct* (*)(const AbstractProduct*)>
class CloneFactory {
public:
    bool Register(const TypeInfo& ti, ProductCreator creator) { return m_map.insert(std::make_pair(ti, creator)).second; }
    bool Unregister(const TypeInfo& ti) { return m_map.erase(ti) == 1; }
    AbstractProduct* CreateObject(const AbstractProduct* model) {
        if (!model) return 0;
        typename std::map<TypeInfo, ProductCreator>::iterator it = m_map.find(TypeInfo(typeid(*model)));
        if (it == m_map.end()) throw std::runtime_error("Unknown Type");
        return (it->second)(model);
    }
private:
    std::map<TypeInfo, ProductCreator> m_map;
};
}
#endif
