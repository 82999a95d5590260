// oracle/ref_tpcc_harness.cpp — TEST INFRASTRUCTURE ONLY (never shipped, never linked into libstage_gpu.so).
//
// The reference's own TPC-C row builders, callable from the tests: BuildStockTuple / BuildOrderLineTuple
// (/root/reference/benchmark/tpcc/tpcc_loader.cpp:1281-1374) compiled UNMODIFIED from where they lie (oracle/Makefile.ref)
// and serialised exactly as the loader stores them: key = Stock::Key().Key() / OrderLine::Key().Key(), payload =
// GetData() (benchmark/tpcc/tpcc_record.h:417-510, 694-772). tests/test_tpcc_rows.py holds the package's seeded
// restatement (stage_indexorganized_b200/tpcc.py) to these bytes: same key bytes, same field offsets, same value
// ranges and the same "one random character repeated" strings (GetRandomAlphaNumericString, tpcc_loader.cpp:849-860).
#include <cstdint>
#include <cstring>

#include "tpcc/tpcc_configuration.h"
#include "tpcc/tpcc_loader.h"
#include "tpcc/tpcc_record.h"

namespace mvstore {
namespace benchmark {
namespace tpcc {
configuration state;        // defined in the driver's main file (benchmark/tpcc/tpcc.cpp:31), which is not linked here
}
}
}

using namespace mvstore::benchmark::tpcc;

extern "C" {

void ref_tpcc_config(int warehouses, int items) {
    state.warehouse_count = warehouses;
    state.item_count = items;
}

int ref_tpcc_stock_sizes(uint32_t *key_size, uint32_t *payload_size) {
    *key_size = sizeof(Stock::StockKey);
    *payload_size = Stock::GetPayloadSize();
    return (int) sizeof(Stock);
}
int ref_tpcc_order_line_sizes(uint32_t *key_size, uint32_t *payload_size) {
    *key_size = sizeof(OrderLine::OrderLineKey);
    *payload_size = OrderLine::GetPayloadSize();
    return (int) sizeof(OrderLine);
}

void ref_tpcc_stock_tuple(int stock_id, int w_id, uint8_t *key_out, uint8_t *payload_out) {
    Stock s = BuildStockTuple(stock_id, w_id);
    auto k = s.Key();
    char *kb = k.Key();
    const char *pb = s.GetData();
    memcpy(key_out, kb, sizeof(Stock::StockKey));
    memcpy(payload_out, pb, Stock::GetPayloadSize());
    delete[] kb;
    delete[] pb;
}

void ref_tpcc_order_line_tuple(int o_id, int d_id, int w_id, int ol_number, int supply_w_id, int new_order, uint8_t *key_out,
                               uint8_t *payload_out) {
    OrderLine ol = BuildOrderLineTuple(o_id, d_id, w_id, ol_number, supply_w_id, new_order != 0);
    auto k = ol.Key();
    char *kb = k.Key();
    const char *pb = ol.GetData();
    memcpy(key_out, kb, sizeof(OrderLine::OrderLineKey));
    memcpy(payload_out, pb, OrderLine::GetPayloadSize());
    delete[] kb;
    delete[] pb;
}

}  // extern "C"
