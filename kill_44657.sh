ssh test@127.0.0.1 "kill -9 2710"
ssh test@127.0.0.1 "kill -9 2711"
kill -9 2109
