// Stand-in for src/core/interfaces/configuration_interface.h:44-59 (same virtuals), plus the in-memory implementation the unit
// tests of the reference use (src/core/receiver/in_memory_configuration.h).
#ifndef B200_STUB_CONFIGURATION_INTERFACE_H
#define B200_STUB_CONFIGURATION_INTERFACE_H
#include <cstdint>
#include <map>
#include <sstream>
#include <string>

class ConfigurationInterface
{
public:
    virtual ~ConfigurationInterface() = default;
    virtual std::string property(std::string property_name, std::string default_value) const = 0;
    virtual bool property(std::string property_name, bool default_value) const = 0;
    virtual int64_t property(std::string property_name, int64_t default_value) const = 0;
    virtual uint64_t property(std::string property_name, uint64_t default_value) const = 0;
    virtual int32_t property(std::string property_name, int32_t default_value) const = 0;
    virtual uint32_t property(std::string property_name, uint32_t default_value) const = 0;
    virtual int16_t property(std::string property_name, int16_t default_value) const = 0;
    virtual uint16_t property(std::string property_name, uint16_t default_value) const = 0;
    virtual float property(std::string property_name, float default_value) const = 0;
    virtual double property(std::string property_name, double default_value) const = 0;
    virtual void set_property(std::string property_name, std::string value) = 0;
};

class InMemoryConfiguration : public ConfigurationInterface
{
public:
    std::string property(std::string n, std::string d) const override
    {
        const auto it = d_map.find(n);
        return it == d_map.end() ? d : it->second;
    }
    bool property(std::string n, bool d) const override
    {
        const auto it = d_map.find(n);
        return it == d_map.end() ? d : (it->second == "true" || it->second == "1");
    }
    int64_t property(std::string n, int64_t d) const override { return get(n, d); }
    uint64_t property(std::string n, uint64_t d) const override { return get(n, d); }
    int32_t property(std::string n, int32_t d) const override { return get(n, d); }
    uint32_t property(std::string n, uint32_t d) const override { return get(n, d); }
    int16_t property(std::string n, int16_t d) const override { return get(n, d); }
    uint16_t property(std::string n, uint16_t d) const override { return get(n, d); }
    float property(std::string n, float d) const override { return get(n, d); }
    double property(std::string n, double d) const override { return get(n, d); }
    void set_property(std::string n, std::string v) override { d_map[n] = v; }

private:
    template <class T>
    T get(const std::string& n, T d) const
    {
        const auto it = d_map.find(n);
        if (it == d_map.end()) return d;
        std::istringstream is(it->second);
        double v = 0.0;  // "4e6" and "4000000.000000" are both written by the reference's tests
        is >> v;
        return is.fail() ? d : static_cast<T>(v);
    }
    std::map<std::string, std::string> d_map;
};
#endif
