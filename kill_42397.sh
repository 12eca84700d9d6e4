ssh test@127.0.0.1 "kill -9 5184"
ssh test@127.0.0.1 "kill -9 5185"
ssh test@127.0.0.1 "kill -9 5186"
ssh test@127.0.0.1 "kill -9 5187"
kill -9 5044
