/* TEST INFRASTRUCTURE — in-process stand-in for ns3-gym's `opengym` module.
 *
 * The reference couples ns-3 to Python through ZMQ + protobuf
 * (ns3-gym/src/opengym/model/opengym_interface.cc:154-286).  Neither library is
 * available in this image, and the transport is exactly the part of the path
 * this project deletes, so the oracle build replaces the module by this header:
 * same class names and the members PRISMA's C++ touches
 * (prisma/ns3/packet-routing-gym.cc:131-213, data-packet-manager.cc:135-206,250-256,
 * sim.cc:538-546,714), with `OpenGymInterface::Notify` asking an in-process
 * policy for the action instead of a socket.  Implementation: ref_harness.cc.
 */
#ifndef PRB_REF_OPENGYM_MODULE_H
#define PRB_REF_OPENGYM_MODULE_H

#include "ns3/object.h"
#include "ns3/ptr.h"
#include "ns3/type-name.h"
#include <stdint.h>
#include <string>
#include <vector>

namespace ns3 {

class OpenGymSpace : public Object
{
public:
  static TypeId GetTypeId ();
  virtual ~OpenGymSpace () {}
};

class OpenGymDiscreteSpace : public OpenGymSpace
{
public:
  static TypeId GetTypeId ();
  OpenGymDiscreteSpace () : m_n (0) {}
  OpenGymDiscreteSpace (int n) : m_n (n) {}
  int GetN () const { return m_n; }
private:
  int m_n;
};

class OpenGymBoxSpace : public OpenGymSpace
{
public:
  static TypeId GetTypeId ();
  OpenGymBoxSpace () : m_low (0), m_high (0) {}
  OpenGymBoxSpace (float low, float high, std::vector<uint32_t> shape, std::string dtype)
    : m_low (low), m_high (high), m_shape (shape), m_dtype (dtype) {}
  float GetLow () const { return m_low; }
  float GetHigh () const { return m_high; }
  std::vector<uint32_t> GetShape () const { return m_shape; }
private:
  float m_low, m_high;
  std::vector<uint32_t> m_shape;
  std::string m_dtype;
};

class OpenGymDataContainer : public Object
{
public:
  static TypeId GetTypeId ();
  virtual ~OpenGymDataContainer () {}
  /* flattened view used by the harness to log observations */
  virtual std::vector<int64_t> AsInt64 () const = 0;
};

class OpenGymDiscreteContainer : public OpenGymDataContainer
{
public:
  static TypeId GetTypeId ();
  OpenGymDiscreteContainer () : m_n (0), m_value (0) {}
  OpenGymDiscreteContainer (uint32_t n) : m_n (n), m_value (0) {}
  bool SetValue (uint32_t value) { m_value = value; return true; }
  uint32_t GetValue () { return m_value; }
  virtual std::vector<int64_t> AsInt64 () const { return std::vector<int64_t> (1, m_value); }
private:
  uint32_t m_n;
  uint32_t m_value;
};

template <typename T = float>
class OpenGymBoxContainer : public OpenGymDataContainer
{
public:
  static TypeId GetTypeId ()
  {
    static TypeId tid = TypeId (("ns3::OpenGymBoxContainer<" + TypeNameGet<T> () + ">").c_str ())
      .SetParent<OpenGymDataContainer> ()
      .SetGroupName ("OpenGym")
      .template AddConstructor<OpenGymBoxContainer<T> > ();
    return tid;
  }
  OpenGymBoxContainer () {}
  OpenGymBoxContainer (std::vector<uint32_t> shape) : m_shape (shape) {}
  bool AddValue (T value) { m_data.push_back (value); return true; }
  T GetValue (uint32_t idx) { return idx < m_data.size () ? m_data[idx] : T (0); }
  std::vector<T> GetData () { return m_data; }
  std::vector<uint32_t> GetShape () { return m_shape; }
  virtual std::vector<int64_t> AsInt64 () const
  {
    std::vector<int64_t> out;
    for (size_t i = 0; i < m_data.size (); i++) out.push_back ((int64_t) m_data[i]);
    return out;
  }
private:
  std::vector<uint32_t> m_shape;
  std::vector<T> m_data;
};

class OpenGymInterface;

class OpenGymEnv : public Object
{
public:
  static TypeId GetTypeId ();
  OpenGymEnv () {}
  virtual ~OpenGymEnv () {}

  virtual Ptr<OpenGymSpace> GetActionSpace () = 0;
  virtual Ptr<OpenGymSpace> GetObservationSpace () = 0;
  virtual bool GetGameOver () = 0;
  virtual Ptr<OpenGymDataContainer> GetObservation () = 0;
  virtual float GetReward () = 0;
  virtual std::string GetExtraInfo () = 0;
  virtual bool ExecuteActions (Ptr<OpenGymDataContainer> action) = 0;

  void SetOpenGymInterface (Ptr<OpenGymInterface> openGymInterface);
  void Notify ();
  void NotifySimulationEnd ();

protected:
  Ptr<OpenGymInterface> m_openGymInterface;
};

class OpenGymInterface : public Object
{
public:
  static TypeId GetTypeId ();
  OpenGymInterface (uint32_t port = 5555);
  virtual ~OpenGymInterface () {}
  void Notify (Ptr<OpenGymEnv> entity);
  void NotifySimulationEnd ();
  uint32_t GetPort () const { return m_port; }
private:
  uint32_t m_port;
  bool m_init;
  int m_index;   /* creation order == overlay node index (sim.cc:528-560) */
  int m_k;       /* size of the action space */
};

} // namespace ns3

#endif
