/* TEST INFRASTRUCTURE — log-component registration the reference gets from libns3-network.
 *
 * In the reference build the vendored ns3-gym/src/network/utils/drop-tail-queue.cc is
 * part of libns3-network and registers the log component "DropTailQueue"; PRISMA's fork
 * (prisma/ns3/drop-tail-queue.h:86) looks that name up at construction while its own .cc
 * registers "DropTailQueueModified" (prisma/ns3/drop-tail-queue.cc:23).  The oracle build
 * leaves the vendored .cc out (its DropTailQueue<Packet> would clash with the fork), so the
 * name is registered here.  No behaviour: logging is compiled out (optimized profile).
 */
#include "ns3/log.h"
namespace ns3 {
NS_LOG_COMPONENT_DEFINE ("DropTailQueue");
}
