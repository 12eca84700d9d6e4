ssh test@127.0.0.1 "kill -9 2835"
ssh test@127.0.0.1 "kill -9 2837"
ssh test@127.0.0.1 "kill -9 2836"
kill -9 2109
