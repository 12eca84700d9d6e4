ssh test@127.0.0.1 "kill -9 3391"
ssh test@127.0.0.1 "kill -9 3392"
ssh test@127.0.0.1 "kill -9 3393"
ssh test@127.0.0.1 "kill -9 3394"
kill -9 2946
