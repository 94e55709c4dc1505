// synerfgine_b200 — present stage entry points + PNG writer (part of the snb_api.cu translation unit).

static FovWarp load_warp(const snb_fov_warp* w) {
	FovWarp r = {0, 0, 0, 1, 0, 0, 0, 0, 0, 1, 0, 1};   // identity: middle segment (y - 0) / 1 everywhere
	if (w) { r.al = w->al; r.bl = w->bl; r.cl = w->cl; r.am = w->am; r.bm = w->bm; r.ar = w->ar; r.br = w->br; r.cr = w->cr;
	         r.switch_left = w->switch_left; r.switch_right = w->switch_right; r.inv_switch_left = w->inv_switch_left; r.inv_switch_right = w->inv_switch_right; }
	return r;
}

extern "C" {

int snb_present(snb_ctx* c, const float* rgba_dev, int in_w, int in_h, int out_w, int out_h, const float* clear_rgb, const snb_fov_warp* warp_x,
                const snb_fov_warp* warp_y, uint8_t* rgb8_dev) {
	if (!c || !rgba_dev || !rgb8_dev || in_w <= 0 || in_h <= 0 || out_w <= 0 || out_h <= 0) return SNB_ERR_ARG;
	SNB_CUDA(c, cudaSetDevice(c->device));
	const float3 clear = clear_rgb ? make_float3(clear_rgb[0], clear_rgb[1], clear_rgb[2]) : make_float3(0, 0, 0);
	const dim3 grid2((out_w + 31) / 32, (out_h + 3) / 4);
	k_present<<<grid2, 128, 0, c->stream>>>(reinterpret_cast<const float4*>(rgba_dev), in_w, in_h, out_w, out_h, clear, load_warp(warp_x), load_warp(warp_y), rgb8_dev);
	SNB_LAUNCHED(c);
	return SNB_OK;
}

int snb_present_host(snb_ctx* c, const float* rgba_dev, int in_w, int in_h, int out_w, int out_h, const float* clear_rgb, const snb_fov_warp* warp_x,
                     const snb_fov_warp* warp_y, uint8_t* rgb8_host) {
	if (!c || !rgb8_host) return SNB_ERR_ARG;
	const size_t bytes = (size_t)out_w * out_h * 3;
	SNB_CUDA(c, cudaSetDevice(c->device));
	SNB_CUDA(c, c->d_present.reserve(bytes));
	const int rc = snb_present(c, rgba_dev, in_w, in_h, out_w, out_h, clear_rgb, warp_x, warp_y, c->d_present.as<uint8_t>());
	if (rc) return rc;
	SNB_CUDA(c, cudaMemcpyAsync(rgb8_host, c->d_present.p, bytes, cudaMemcpyDeviceToHost, c->stream));
	SNB_CUDA(c, cudaStreamSynchronize(c->stream));
	return SNB_OK;
}

// Minimal PNG (8-bit RGB, filter 0, zlib stream of stored deflate blocks): decodes to exactly the bytes given. The reference
// writes through stb_image_write (display.cu:318-319), whose own deflate makes the FILE bytes implementation-specific; the
// decoded pixels are what is comparable.
int snb_write_png(const char* path, const uint8_t* rgb8_host, int width, int height) {
	if (!path || !rgb8_host || width <= 0 || height <= 0) return SNB_ERR_ARG;
	static uint32_t crc_table[256];
	static bool have_table = false;
	if (!have_table) {
		for (uint32_t n = 0; n < 256; ++n) { uint32_t v = n; for (int k = 0; k < 8; ++k) v = (v & 1u) ? 0xEDB88320u ^ (v >> 1) : v >> 1; crc_table[n] = v; }
		have_table = true;
	}
	auto crc = [&](const uint8_t* p, size_t n, uint32_t v) { for (size_t i = 0; i < n; ++i) v = crc_table[(v ^ p[i]) & 0xFFu] ^ (v >> 8); return v; };
	auto be32 = [](uint8_t* p, uint32_t v) { p[0] = (uint8_t)(v >> 24); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t)v; };
	FILE* f = std::fopen(path, "wb");
	if (!f) return SNB_ERR_IO;
	auto chunk = [&](const char* type, const std::vector<uint8_t>& data) {
		uint8_t hdr[8]; be32(hdr, (uint32_t)data.size()); std::memcpy(hdr + 4, type, 4);
		uint32_t v = crc(hdr + 4, 4, 0xFFFFFFFFu); v = crc(data.data(), data.size(), v) ^ 0xFFFFFFFFu;
		uint8_t tail[4]; be32(tail, v);
		return std::fwrite(hdr, 1, 8, f) == 8 && (data.empty() || std::fwrite(data.data(), 1, data.size(), f) == data.size()) && std::fwrite(tail, 1, 4, f) == 4;
	};
	static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
	bool ok = std::fwrite(sig, 1, 8, f) == 8;
	std::vector<uint8_t> ihdr(13);
	be32(ihdr.data(), (uint32_t)width); be32(ihdr.data() + 4, (uint32_t)height);
	ihdr[8] = 8; ihdr[9] = 2; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;                 // 8 bit, colour type 2 (RGB), deflate, adaptive filters, no interlace
	ok = ok && chunk("IHDR", ihdr);
	const size_t row = (size_t)width * 3 + 1, raw_bytes = row * (size_t)height;
	std::vector<uint8_t> raw(raw_bytes);
	for (int y = 0; y < height; ++y) { raw[y * row] = 0; std::memcpy(&raw[y * row + 1], rgb8_host + (size_t)y * width * 3, (size_t)width * 3); }
	std::vector<uint8_t> z; z.reserve(raw_bytes + raw_bytes / 65535 * 5 + 16);
	z.push_back(0x78); z.push_back(0x01);
	uint32_t s1 = 1, s2 = 0;
	for (size_t off = 0; off < raw_bytes || off == 0; ) {
		const size_t n = std::min<size_t>(65535, raw_bytes - off);
		z.push_back(off + n >= raw_bytes ? 1 : 0);
		z.push_back((uint8_t)(n & 0xFF)); z.push_back((uint8_t)(n >> 8)); z.push_back((uint8_t)(~n & 0xFF)); z.push_back((uint8_t)((~n >> 8) & 0xFF));
		z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
		for (size_t i = 0; i < n; ++i) { s1 = (s1 + raw[off + i]) % 65521u; s2 = (s2 + s1) % 65521u; }
		off += n;
		if (n == 0) break;
	}
	uint8_t ad[4]; be32(ad, (s2 << 16) | s1); z.insert(z.end(), ad, ad + 4);
	ok = ok && chunk("IDAT", z) && chunk("IEND", std::vector<uint8_t>());
	ok = (std::fclose(f) == 0) && ok;
	return ok ? SNB_OK : SNB_ERR_IO;
}

} // extern "C"
